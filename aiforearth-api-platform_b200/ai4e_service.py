"""`ai4e_service` - drop-in for the container host shell of the AI for Earth API Platform.

Same module name, classes, decorator signatures, URL rules, responses and task-status
transitions as the reference (/root/reference/APIs/1.0/base-py/ai4e_service.py:30-212), so an
existing API project imports it unchanged:

    from ai4e_service import APIService
    ai4e = APIService(app, log)

    @ai4e.api_async_func(api_path='/classify', methods=['POST'],
                         request_processing_function=read_image, maximum_concurrent_requests=512,
                         content_types=['application/octet-stream'], content_max_length=10_000_000,
                         trace_name='post:classify')
    def classify(*args, **kwargs):
        taskId, image = kwargs['taskId'], kwargs['image']
        ai4e.api_task_manager.UpdateTaskStatus(taskId, 'running')
        idx, prob = engine.classify(image)                 # <- the B200 engine (ai4e_b200.engine)
        ai4e.api_task_manager.CompleteTask(taskId, 'completed - ' + json.dumps(idx.tolist()))

Contract kept (reference line numbers):
  * `APIService(flask_app, logger)` exposes .app .api .log .tracer .api_prefix
    .api_task_manager .func_properties .func_request_counts .is_terminating (:44-56);
  * `api_sync_func` / `api_async_func` / `api_func` take the same arguments (:72,103-109) and
    register `{API_PREFIX}{api_path}` under the view name `api_<path without slashes>` (:98-100);
  * the task function is called as `func(*args, **kwargs)` with kwargs `func`, `api_path`,
    whatever `request_processing_function(request)` returned, URL values, and for async
    endpoints `taskId` and `request` (:80-91,181,203);
  * a sync endpoint answers with the function's return value; an async endpoint answers
    `'TaskId: <id>'` at once and runs the function on another thread (:89-96);
  * errors: an HTTPException raised by the function becomes the response, anything else a 500;
    async failures mark the task `FailTask(taskId, 'Task failed - try again')` (:185-211);
  * admission: terminating or over `maximum_concurrent_requests` -> 503, wrong content type ->
    401, body too large -> 413 (:116-133); `GET {prefix}/` -> 'Health check OK' (:60,68-70);
    `GET {prefix}/task/<id>` -> the task's status (:30-41,63).

Built differently underneath:
  * async tasks run on reusable worker threads instead of one new `threading.Thread` per task
    (:180-183); many workers blocked in `engine.classify` are what fills the engine's
    micro-batches.  A task function may instead `return engine.classify_later(image, then=...)`:
    the worker is released at once, the engine's single reaper thread runs `then` (which calls
    CompleteTask) when the micro-batch has finished, and the shell keeps the request counted and
    maps a failure to FailTask until then - no thread is blocked per task.  Such a non-blocking
    function may also be registered with `inline=True` (`api_async_func(..., inline=True)`): it
    then runs on the request thread and a task needs no thread hand-off on its way in;
  * in-flight counters are guarded by a lock (:148-156 is unsynchronised);
  * `kwargs['request']` is a snapshot that stays readable on the worker thread (the reference
    passes the context-bound proxy, :181);
  * admission looks the path up with API_PREFIX stripped, so the checks also fire when a prefix
    is set (:74-76 registers `api_path`, :122 looks up `request.path`), and
    `maximum_concurrent_requests=None` means "no limit" rather than a TypeError;
  * AI4E_BUSY_STATUS=429 makes a busy service answer the status the platform's queue pusher
    retries on (ProcessManager/BackendQueueProcessor/BackendQueueProcessor.cs:54-64); the
    default stays 503;
  * the task resource reads the status dict (the reference indexes it like a tuple, :35-39).
"""
import contextlib
import json
import logging
import queue
import signal
import sys
import threading
from os import getenv

try:
    from flask import abort, request
    from flask_restful import Api, Resource
except ImportError:  # neither is in the build image; same surface on Werkzeug
    from ai4e_b200.compat.miniflask import abort, request
    from ai4e_b200.compat.minirestful import Api, Resource
from werkzeug.datastructures import Headers
from werkzeug.exceptions import HTTPException

from task_management.api_task import TaskManager

try:
    from ai4e_app_insights_wrapper import AI4EAppInsights
except ImportError:
    class AI4EAppInsights:  # keeps isinstance() meaningful without the Azure wrapper installed
        pass

_log = logging.getLogger("ai4e.service")

disable_request_metric = getenv('DISABLE_CURRENT_REQUEST_METRIC', 'True')
current_processing_upsert_url = getenv('CURRENT_PROCESSING_UPSERT_URI')
service_cluster = getenv('SERVICE_CLUSTER', 'undefined')

MAX_REQUESTS_KEY_NAME = 'max_requests'
CONTENT_TYPE_KEY_NAME = 'content_types'
CONTENT_MAX_KEY_NAME = 'content_max_length'
APP_INSIGHTS_REQUESTS_KEY_NAME = 'REJECTED_STATE'

BUSY_MESSAGE = {'message': 'Service is busy, please try again later.'}
TASK_FAILED_STATUS = 'Task failed - try again'


def _request_metric_enabled():
    off = str(disable_request_metric).strip().lower() not in ('false', '0', 'no')
    return not off and bool(current_processing_upsert_url)


class RequestSnapshot:
    """What a task function reads from a request, copied while the request context is alive."""

    def __init__(self, live):
        self.method, self.path = live.method, live.path
        self.url = getattr(live, 'url', live.path)
        self.headers = Headers(list(live.headers.items()))  # case-insensitive, like the live object
        self.args = dict(live.args.items()) if getattr(live, 'args', None) is not None else {}
        self.content_type, self.content_length = live.content_type, live.content_length
        try:
            self.data = live.get_data()
        except Exception:  # noqa: BLE001 - an already consumed stream is not an error here
            self.data = b''

    def get_data(self, as_text=False):
        return self.data.decode('utf-8') if as_text else self.data

    def get_json(self, silent=False):
        try:
            return json.loads(self.data)
        except ValueError:
            if silent:
                return None
            raise

    json = property(lambda self: self.get_json(silent=True))


class _TaskWorkers:
    """Daemon worker threads that grow on demand up to a cap, then queue."""

    def __init__(self, cap):
        self._cap, self._jobs = cap, queue.SimpleQueue()
        self._guard = threading.Lock()
        self._idle = self._alive = self._pending = 0

    def run(self, fn, args, kwargs):
        # a job starts at once whenever a thread may still be created: an idle worker counts only if
        # no queued job has already claimed it (two requests arriving back to back must not share one)
        with self._guard:
            self._pending += 1
            grow = self._idle < self._pending and self._alive < self._cap
            self._alive += grow
        self._jobs.put((fn, args, kwargs))
        if grow:
            threading.Thread(target=self._loop, daemon=True, name='ai4e-task').start()

    def queued(self):
        """jobs accepted but not yet picked up by a worker (only non-zero at the thread cap)."""
        with self._guard:
            return max(0, self._pending - self._idle)

    def _loop(self):
        while True:
            with self._guard:
                self._idle += 1
            fn, args, kwargs = self._jobs.get()
            with self._guard:
                self._idle -= 1
                self._pending -= 1
            try:
                fn(*args, **kwargs)
            except BaseException:  # noqa: BLE001 - already logged and the task failed by the wrapper
                pass


class _CurrentTaskManager:
    """forwards to whatever `APIService.api_task_manager` is now (tests and in-process platforms
    swap it after construction; the reference captures the object once, :63)."""

    def __init__(self, service):
        self._service = service

    def GetTaskStatus(self, task_id):
        return self._service.api_task_manager.GetTaskStatus(task_id)


class Task(Resource):
    def __init__(self, **kwargs):
        self.task_mgr = kwargs['task_manager']

    def get(self, id):
        status = self.task_mgr.GetTaskStatus(str(id))
        return {'TaskId': id, 'Status': status.get('Status'), 'Timestamp': status.get('Timestamp'),
                'Endpoint': status.get('Endpoint', 'uri')}


class APIService():
    def __init__(self, flask_app, logger):
        self.app, self.log = flask_app, logger
        self.api = Api(self.app)
        self.api_prefix = getenv('API_PREFIX') or ''
        self.is_terminating = False
        self.func_properties, self.func_request_counts = {}, {}
        # an AI4EAppInsights logger carries no tracer; anything else may
        self.tracer = None if isinstance(logger, AI4EAppInsights) else getattr(logger, 'tracer', None)
        self.api_task_manager = TaskManager()

        self._counts_guard = threading.Lock()
        self._workers = _TaskWorkers(int(getenv('AI4E_B200_MAX_WORKERS', '1024')))
        self._busy_status = int(getenv('AI4E_BUSY_STATUS', '503'))
        with contextlib.suppress(ValueError):  # signal handlers can only be set on the main thread
            signal.signal(signal.SIGINT, self.initialize_term)

        self.app.add_url_rule(self.api_prefix + '/', view_func=self.health_check, methods=['GET'])
        self.api.add_resource(Task, self.api_prefix + '/task/<id>',
                              resource_class_kwargs={'task_manager': _CurrentTaskManager(self)})
        self.app.before_request(self.before_request)
        _log.info("host shell ready, prefix %r", self.api_prefix)

    # ------------------------------------------------------------------ endpoints
    def health_check(self):
        return 'Health check OK'

    # ------------------------------------------------------------------ autoscaling signals
    def attach_engine(self, engine, route=True):
        """the B200 engine (or pool) whose queue depth and latency `metrics()` should report; also
        serves `metrics()` as JSON at `{prefix}/metrics` (call at start-up, before the first request;
        the reference has no such route, so it only exists once this is called)."""
        self._engine = engine
        if route and not getattr(self, '_metrics_routed', False):
            self.app.add_url_rule(self.api_prefix + '/metrics', view_func=lambda: json.dumps(self.metrics()),
                                  methods=['GET'], endpoint='ai4e_metrics')
            self._metrics_routed = True

    def metrics(self):
        """In-flight requests per endpoint under the name the platform's HPA consumes,
        `CURRENT_REQUESTS/{cluster}{prefix}{path}` (ProcessManager/RequestReporter/
        CurrentProcessingUpsert.cs:102-104, APIs/Charts/templates/appinsights-metric.yaml:7), plus,
        when an engine is attached, what the reference cannot see: queue depth, batch sizes and
        task latency of the replica(s)."""
        with self._counts_guard:
            counts = dict(self.func_request_counts)
        out = {'CURRENT_REQUESTS/%s%s%s' % (service_cluster, self.api_prefix, path): n for path, n in counts.items()}
        engine = getattr(self, '_engine', None)
        if engine is not None:
            stats = engine.stats()
            for i, st in enumerate(stats if isinstance(stats, list) else [stats]):
                for key in ('queue_depth', 'images', 'batches', 'lat_ms_p50', 'lat_ms_p99', 'healthy'):
                    out['ENGINE/%s/replica%d/%s' % (service_cluster, i, key)] = st[key]
        return out

    def api_func(self, is_async, api_path, methods, request_processing_function, maximum_concurrent_requests,
                 content_types=None, content_max_length=None, trace_name=None, *args, **kwargs):
        # `inline=True` (through the reference's unused **kwargs, so the signature stays as it is): the task
        # function of an async endpoint runs on the request thread itself.  Only for functions that do not
        # block - they hand their work to the engine with `classify_later` and return the Pending - and
        # then a task costs no thread hand-off at all on its way in.
        inline = bool(kwargs.pop('inline', False)) and is_async

        def register(func):
            if api_path not in self.func_properties:
                self.func_properties[api_path] = {MAX_REQUESTS_KEY_NAME: maximum_concurrent_requests,
                                                  CONTENT_TYPE_KEY_NAME: content_types,
                                                  CONTENT_MAX_KEY_NAME: content_max_length}
                self.func_request_counts[api_path] = 0

            def view(*view_args, **url_values):
                call_kwargs = {'func': func, 'api_path': api_path, **url_values}
                if request_processing_function:
                    call_kwargs.update(request_processing_function(request))
                if not is_async:
                    return self.wrap_sync_endpoint(trace_name, *view_args, **call_kwargs)
                task = self.api_task_manager.AddTask(request)
                call_kwargs['taskId'] = str(task['TaskId'])
                if inline:
                    call_kwargs['request'] = self._detached_request()
                    with self._span(trace_name, call_kwargs):
                        self._execute_func_with_counter(True, *view_args, **call_kwargs)
                else:
                    self.wrap_async_endpoint(trace_name, *view_args, **call_kwargs)
                return 'TaskId: ' + call_kwargs['taskId']

            view.__name__ = 'api_' + api_path.replace('/', '')
            view.__doc__ = func.__doc__
            view.__wrapped__ = func
            self.app.add_url_rule(self.api_prefix + api_path, view_func=view, methods=methods,
                                  provide_automatic_options=True)
            _log.info("url rule %s%s -> %s", self.api_prefix, api_path, view.__name__)
            return func
        return register

    def api_async_func(self, api_path, methods, request_processing_function=None, maximum_concurrent_requests=None,
                       content_types=None, content_max_length=None, trace_name=None, *args, **kwargs):
        return self.api_func(True, api_path, methods, request_processing_function, maximum_concurrent_requests,
                             content_types, content_max_length, trace_name, *args, **kwargs)

    def api_sync_func(self, api_path, methods, request_processing_function=None, maximum_concurrent_requests=None,
                      content_types=None, content_max_length=None, trace_name=None, *args, **kwargs):
        return self.api_func(False, api_path, methods, request_processing_function, maximum_concurrent_requests,
                             content_types, content_max_length, trace_name, *args, **kwargs)

    # ------------------------------------------------------------------ admission
    def initialize_term(self, signum, frame):
        _log.warning("signal %s: no longer accepting requests", signum)
        self.is_terminating = True

    def _registered_path(self, path):
        """`request.path` -> the `api_path` it was registered under, or None."""
        if path in self.func_properties:
            return path
        tail = path[len(self.api_prefix):] if self.api_prefix and path.startswith(self.api_prefix) else None
        return tail if tail in self.func_properties else None

    def before_request(self):
        if self.is_terminating:
            abort(503, BUSY_MESSAGE)
        api_path = self._registered_path(request.path)
        if api_path is None:
            return
        rules = self.func_properties[api_path]
        limit, types, max_len = rules[MAX_REQUESTS_KEY_NAME], rules[CONTENT_TYPE_KEY_NAME], rules[CONTENT_MAX_KEY_NAME]
        if limit is not None and self.func_request_counts[api_path] + 1 > limit:
            abort(self._busy_status, BUSY_MESSAGE)
        if types and request.content_type not in types:
            abort(401, {'message': 'Content-type must be ' + str(types)})
        size = request.content_length
        if max_len and size is not None and size > max_len:
            abort(413, {'message': 'Request content too large (%s). Must be smaller than: %s' % (size, max_len)})

    # ------------------------------------------------------------------ in-flight accounting
    def update_processing_count(self, api_path, increment_by, decrement_by):
        import requests
        payload = {'ApiPath': self.api_prefix + api_path, 'ServiceCluster': service_cluster,
                   'IncrementBy': increment_by, 'DecrementBy': decrement_by}
        try:
            reply = requests.post(current_processing_upsert_url, data=json.dumps(payload), timeout=5)
            if reply.status_code != 200:
                _log.warning("request metric answered %s", reply.status_code)
        except Exception as err:  # noqa: BLE001 - a metric must never fail a request
            _log.warning("request metric failed: %s", err)

    def _bump(self, api_path, delta):
        with self._counts_guard:
            self.func_request_counts[api_path] += delta
        if _request_metric_enabled():
            self.update_processing_count(api_path, increment_by=max(delta, 0), decrement_by=max(-delta, 0))

    def increment_requests(self, api_path):
        self._bump(api_path, +1)

    def decrement_requests(self, api_path):
        self._bump(api_path, -1)

    # ------------------------------------------------------------------ dispatch
    def _span(self, trace_name, call_kwargs):
        if not self.tracer:
            return contextlib.nullcontext()
        return self.tracer.span(name=trace_name or call_kwargs['api_path'])

    def wrap_sync_endpoint(self, trace_name=None, *args, **kwargs):
        with self._span(trace_name, kwargs):
            return self._execute_func_with_counter(False, *args, **kwargs)

    def wrap_async_endpoint(self, trace_name=None, *args, **kwargs):
        with self._span(trace_name, kwargs):
            self._create_and_execute_thread(*args, **kwargs)

    @staticmethod
    def _detached_request():
        """the request as the task function may read it on another thread.  The real request object
        behind the context-bound proxy stays valid once its body is cached, and keeps the complete
        API (headers, args, json, files); only a foreign request type falls back to a copy."""
        live = request
        unwrap = getattr(live, '_get_current_object', None)
        if unwrap is None:
            return RequestSnapshot(live)
        obj = unwrap()
        try:
            obj.get_data(cache=True)  # the server may close the input stream once the view has answered
        except Exception:  # noqa: BLE001 - not a Werkzeug request after all
            return RequestSnapshot(live)
        return obj

    def _create_and_execute_thread(self, *args, **kwargs):
        kwargs['request'] = self._detached_request()
        self._workers.run(self._execute_func_with_counter, args, kwargs)

    def _log_and_fail_exeception(self, is_async, **kwargs):
        task_id = kwargs.get('taskId')
        if not task_id:
            self.log.log_exception(sys.exc_info()[0])
            return
        self.log.log_exception(sys.exc_info()[0], task_id)
        if is_async:
            self.api_task_manager.FailTask(task_id, TASK_FAILED_STATUS)

    def _execute_func_with_counter(self, is_async=True, *args, **kwargs):
        func, api_path = kwargs['func'], kwargs['api_path']
        self.increment_requests(api_path)
        deferred = False
        try:
            result = func(*args, **kwargs)
            if is_async and hasattr(result, 'add_done_callback') and hasattr(result, 'error'):
                # the function handed its image to the engine with `classify_later` and returned the
                # Pending: the task is still in flight.  Its worker thread is free now; the request
                # stays counted, and a failure is reported, when the engine's reaper resolves it.
                deferred = True
                result.add_done_callback(lambda pending: self._finish_deferred(pending, kwargs))
            return result
        except HTTPException as http_error:  # the function chose its own HTTP answer
            self._log_and_fail_exeception(is_async, **kwargs)
            return http_error
        except Exception:  # noqa: BLE001 - upstream catches everything here too
            self._log_and_fail_exeception(is_async, **kwargs)
            if not is_async:  # on a worker thread there is no response left to turn into a 500
                abort(500)
        finally:
            if not deferred:
                self.decrement_requests(api_path)

    def _finish_deferred(self, pending, kwargs):
        """runs on the engine's reaper thread when a deferred task's result (and its `then`) is through."""
        try:
            if pending.error is not None:
                task_id = kwargs.get('taskId')
                try:
                    self.log.log_exception(type(pending.error), task_id)
                except Exception:  # noqa: BLE001 - a logger must not keep the task from being failed
                    pass
                if task_id:
                    self.api_task_manager.FailTask(task_id, TASK_FAILED_STATUS)
        finally:
            self.decrement_requests(kwargs['api_path'])
