"""One scripted conversation with an `APIService`, used twice: against the UNMODIFIED reference
class (tests/golden/make_host_shell_trace.py, run where /root/reference exists) and against
this repo's `ai4e_service.APIService` (tests/test_host_shell.py).  It records what a client and
a task function can observe: URL rules, status codes, bodies, the kwargs that reach `func`,
the task-manager call sequence and the in-flight counters."""
import json
import threading
import time

from werkzeug.exceptions import NotFound


class RecordingTaskManager:
    """In-memory stand-in with the TaskManager method names; records every call."""

    def __init__(self):
        self.calls = []
        self.lock = threading.Lock()
        self.done = threading.Event()

    def _rec(self, *c):
        with self.lock:
            self.calls.append(list(c))

    def AddTask(self, request):
        has = bool(request) and 'taskId' in request.headers
        self._rec('AddTask', has)
        return {'TaskId': request.headers.get('taskId') if has else '-1', 'Status': 'created' if has else 'error'}

    def UpdateTaskStatus(self, taskId, status):
        self._rec('UpdateTaskStatus', taskId, status)
        return {'TaskId': taskId, 'Status': status}

    def CompleteTask(self, taskId, status):
        self._rec('CompleteTask', taskId, status)
        self.done.set()
        return {'TaskId': taskId, 'Status': status}

    def FailTask(self, taskId, status):
        self._rec('FailTask', taskId, status)
        self.done.set()

    def GetTaskStatus(self, taskId):
        self._rec('GetTaskStatus', taskId)
        return {'TaskId': taskId, 'Status': 'created', 'Timestamp': 't', 'Endpoint': 'http://h/v1/x'}


class Logger:
    tracer = None

    def __init__(self):
        self.exceptions = 0

    def log_exception(self, *a, **k):
        self.exceptions += 1


def run_scenario(APIService, Flask):
    app = Flask('scenario')
    log = Logger()
    svc = APIService(app, log)
    tm = RecordingTaskManager()
    svc.api_task_manager = tm
    seen = {}

    def read_body(request):
        return {'image_bytes': request.get_data()}

    @svc.api_sync_func(api_path='/echo', methods=['POST'], request_processing_function=read_body,
                       maximum_concurrent_requests=4, content_types=['application/octet-stream'],
                       content_max_length=16, trace_name='post:echo')
    def echo(*args, **kwargs):
        seen['sync_kwargs'] = sorted(kwargs)
        seen['sync_counter_inside'] = svc.func_request_counts['/echo']
        return 'sync:%d' % len(kwargs['image_bytes'])

    @svc.api_sync_func(api_path='/item/<name>', methods=['GET'], maximum_concurrent_requests=4)
    def item(*args, **kwargs):
        seen['url_kwargs'] = sorted(kwargs)
        return 'item:' + kwargs['name']

    @svc.api_sync_func(api_path='/missing', methods=['GET'], maximum_concurrent_requests=4)
    def missing(*args, **kwargs):
        raise NotFound()

    @svc.api_sync_func(api_path='/boom', methods=['GET'], maximum_concurrent_requests=4)
    def boom(*args, **kwargs):
        raise ValueError('boom')

    @svc.api_sync_func(api_path='/full', methods=['GET'], maximum_concurrent_requests=0)
    def full(*args, **kwargs):
        return 'never'

    @svc.api_async_func(api_path='/classify', methods=['POST'], request_processing_function=read_body,
                        maximum_concurrent_requests=8, content_types=['application/octet-stream'])
    def classify(*args, **kwargs):
        seen['async_kwargs'] = sorted(kwargs)
        seen['async_thread_is_request_thread'] = threading.current_thread() is seen.get('request_thread')
        svc.api_task_manager.UpdateTaskStatus(kwargs['taskId'], 'running')
        svc.api_task_manager.CompleteTask(kwargs['taskId'], 'completed - %d bytes' % len(kwargs['image_bytes']))

    @svc.api_async_func(api_path='/classify_fail', methods=['POST'], maximum_concurrent_requests=8)
    def classify_fail(*args, **kwargs):
        raise ValueError('bad image')

    out = {'rules': sorted([r.rule, sorted(m for m in r.methods if m not in ('HEAD',)), r.endpoint]
                           for r in app.url_map.iter_rules())}
    c = app.test_client()
    octet = {'Content-Type': 'application/octet-stream'}

    def hit(name, method, path, **kw):
        r = getattr(c, method)(path, **kw)
        out[name] = {'status': r.status_code, 'body': r.get_data(as_text=True) if r.status_code < 300 else None}

    seen['request_thread'] = threading.current_thread()
    hit('health', 'get', '/')
    hit('sync_ok', 'post', '/echo', data=b'abc', headers=octet)
    hit('sync_bad_type', 'post', '/echo', data=b'abc', headers={'Content-Type': 'text/plain'})
    hit('sync_too_large', 'post', '/echo', data=b'x' * 32, headers=octet)
    hit('sync_url_value', 'get', '/item/fox')
    hit('sync_http_exception', 'get', '/missing')
    hit('sync_exception', 'get', '/boom')
    hit('sync_busy', 'get', '/full')
    hit('unknown_path', 'get', '/nope')

    tm.done.clear()
    hit('async_ok', 'post', '/classify', data=b'12345', headers={**octet, 'taskId': '42'})
    assert tm.done.wait(5)
    tm.done.clear()
    hit('async_fail', 'post', '/classify_fail', data=b'', headers={'taskId': '43'})
    assert tm.done.wait(5)
    for _ in range(200):  # the worker decrements its counter after the task function returns
        if not any(svc.func_request_counts.values()):
            break
        time.sleep(0.01)

    svc.is_terminating = True
    hit('terminating', 'get', '/')
    svc.is_terminating = False

    seen.pop('request_thread')
    out['seen'] = seen
    out['task_calls'] = tm.calls
    out['counters_after'] = dict(sorted(svc.func_request_counts.items()))
    out['logged_exceptions'] = log.exceptions
    out['attributes'] = sorted(a for a in ('app', 'api', 'log', 'tracer', 'api_prefix', 'api_task_manager',
                                           'func_properties', 'func_request_counts', 'is_terminating')
                               if hasattr(svc, a))
    out['func_properties_echo'] = svc.func_properties['/echo']
    return json.loads(json.dumps(out))
