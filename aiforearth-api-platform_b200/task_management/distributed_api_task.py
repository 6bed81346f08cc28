"""Client of the platform's two cache-connector functions - same class, methods, payloads and
sentinel results as the reference's `DistributedApiTaskManager`
(/root/reference/APIs/1.0/Common/task_management/distributed_api_task.py:12-116).

Differences, all on the transport and none on the contract:
  * the reference opens a new aiohttp session and a new event loop for every call (:18,37,
    79-80,103-104); this client keeps one pooled `requests.Session` per thread;
  * when no CACHE_CONNECTOR_*_URI is configured, or AI4E_TASK_STORE=local, calls go to the
    in-process `LocalTaskStore` (the control plane is out of scope and is exercised in-process);
  * the error branches that raise NameError/AttributeError upstream (:50,94) return the
    documented sentinel dicts instead;
  * opt-in, AI4E_TASK_COALESCE=1 (or `coalesce=True`): the Endpoint a task was created with is
    remembered from the first read, so a status change is ONE upsert instead of the reference's
    read-modify-write (:30 + :42) - 3 round trips per task (AddTask's read, running, completed)
    instead of 5;
  * opt-in, AI4E_TASK_WRITE_BEHIND=1 (or `write_behind=True`, implies coalescing): status writes are
    queued to one writer thread and sent in order; the caller gets the task dict it wrote at once
    and never waits for the platform.  `flush()` drains the queue (call it before exiting).
"""
from __future__ import annotations

import datetime
import time
import json
import queue
import threading
from collections import OrderedDict
from os import getenv
from urllib.parse import urlparse

from task_management.local_task_store import LocalTaskStore, default_store

_NOT_FOUND = '{{"TaskId": "{0}", "Status": "not found"}}'
_ERROR = '{"TaskId": "-1", "Status": "error"}'


def _flag(name):
    return getenv(name, '').strip().lower() in ('1', 'true', 'yes')


class DistributedApiTaskManager:
    _ENDPOINT_CACHE = 65536  # tasks whose Endpoint is remembered (coalescing); oldest dropped first

    def __init__(self, store: LocalTaskStore | None = None, coalesce: bool | None = None,
                 write_behind: bool | None = None):
        self.cache_connector_upsert_url = getenv('CACHE_CONNECTOR_UPSERT_URI')
        self.cache_connector_get_url = getenv('CACHE_CONNECTOR_GET_URI')
        use_local = store is not None or getenv('AI4E_TASK_STORE', '').lower() == 'local' or not (
            self.cache_connector_upsert_url and self.cache_connector_get_url)
        self.store = (store or default_store()) if use_local else None
        self._tls = threading.local()
        self.write_behind = _flag('AI4E_TASK_WRITE_BEHIND') if write_behind is None else bool(write_behind)
        self.coalesce = self.write_behind or (_flag('AI4E_TASK_COALESCE') if coalesce is None else bool(coalesce))
        self._endpoints: OrderedDict[str, str] = OrderedDict()
        self._endpoints_guard = threading.Lock()
        self._writes: queue.Queue | None = None
        self._writer = None
        self._writer_guard = threading.Lock()

    # ---- coalescing: remember each task's Endpoint so that a status change needs no read
    def _remember(self, task_id, status):
        if not self.coalesce or not isinstance(status, dict):
            return
        endpoint = status.get('Endpoint')
        if not endpoint:
            return
        with self._endpoints_guard:
            self._endpoints[str(task_id)] = endpoint
            self._endpoints.move_to_end(str(task_id))
            while len(self._endpoints) > self._ENDPOINT_CACHE:
                self._endpoints.popitem(last=False)

    def _known_endpoint(self, task_id, forget=False):
        if not self.coalesce:
            return None
        with self._endpoints_guard:
            return self._endpoints.pop(str(task_id), None) if forget else self._endpoints.get(str(task_id))

    # ---- write-behind: one writer thread keeps the order of the status writes
    def _enqueue(self, payload):
        with self._writer_guard:
            if self._writer is None:
                self._writes = queue.Queue()
                self._writer = threading.Thread(target=self._drain, daemon=True, name='ai4e-task-writer')
                self._writer.start()
        self._writes.put(payload)

    def _drain(self):
        while True:
            payload = self._writes.get()
            try:
                code, _ = self._post(payload)
                if code != 200:
                    print("status code: " + str(code))
            finally:
                self._writes.task_done()

    def flush(self):
        """block until every queued status write has been sent (write-behind only)."""
        if self._writes is not None:
            self._writes.join()

    # ---- transport
    def _session(self):
        s = getattr(self._tls, "session", None)
        if s is None:
            import requests
            s = requests.Session()
            self._tls.session = s
        return s

    def _post(self, payload):
        """-> (status code, stored task dict or None)"""
        if self.store is not None:  # in-process store: the dict itself, no JSON round trip
            return self.store.upsert_task(payload) if payload is not None else (400, None)
        body = json.dumps(payload) if payload is not None else None
        try:
            r = self._session().post(self.cache_connector_upsert_url, data=body,
                                     headers={"Content-Type": "application/json"}, timeout=30)
            return r.status_code, (json.loads(r.text) if r.status_code == 200 else None)
        except Exception as e:  # noqa: BLE001
            print("task store POST failed: " + str(e))
            return 599, None

    def _get(self, task_id):
        """-> (status code, task dict or None)"""
        if self.store is not None:
            t = self.store.get_task(task_id)
            return (200, t) if t is not None else (204, None)
        try:
            r = self._session().get(self.cache_connector_get_url, params={'taskId': task_id}, timeout=30)
            return r.status_code, (json.loads(r.text) if r.status_code == 200 else None)
        except Exception as e:  # noqa: BLE001
            print("task store GET failed: " + str(e))
            return 599, None

    _now_cache = (0, "")

    @classmethod
    def _now(cls):
        """local "YYYY-MM-DD HH:MM:SS" (the reference's format, :33), formatted once per second"""
        sec = int(time.time())
        if cls._now_cache[0] != sec:
            cls._now_cache = (sec, datetime.datetime.fromtimestamp(sec).strftime("%Y-%m-%d %H:%M:%S"))
        return cls._now_cache[1]

    # ---- the reference's public methods
    def AddTask(self):
        code, task = self._post(None)  # an empty body: the platform answers 400 -> the error sentinel
        return task if code == 200 else json.loads(_ERROR)

    def _UpdateTaskStatus(self, taskId, status, backendStatus):
        final = backendStatus in ('completed', 'failed')
        endpoint = self._known_endpoint(taskId, forget=final)
        if endpoint is None:  # the reference's path: read, then write (:30-38)
            old_stat = self.GetTaskStatus(taskId)
            endpoint = 'http://localhost'
            if not old_stat.get('Endpoint'):
                print("Cannot find task status. Creating")
            else:
                endpoint = old_stat['Endpoint']
            if final:
                self._known_endpoint(taskId, forget=True)
        payload = {'TaskId': taskId, 'Timestamp': self._now(), 'Status': status,
                   'BackendStatus': backendStatus, 'Endpoint': endpoint, 'PublishToGrid': False}
        if self.write_behind:
            self._enqueue(payload)
            return dict(payload)
        code, task = self._post(payload)
        if code != 200:
            print("status code: " + str(code))
            return json.loads(_NOT_FOUND.format(taskId))
        return task

    def CompleteTask(self, taskId, status):
        return self._UpdateTaskStatus(taskId, status, 'completed')

    def UpdateTaskStatus(self, taskId, status):
        return self._UpdateTaskStatus(taskId, status, 'running')

    def FailTask(self, taskId, status):
        return self._UpdateTaskStatus(taskId, status, 'failed')

    def AddPipelineTask(self, taskId, organization_moniker, version, api_name, body):
        old_stat = self.GetTaskStatus(taskId)
        if old_stat['Status'] == "not found":
            print("Cannot find task status.")
            return json.loads(_ERROR)
        parsed_endpoint = urlparse(old_stat['Endpoint'])
        path = '{}/{}/{}'.format(version, organization_moniker, api_name)
        next_endpoint = '{}://{}/{}'.format(parsed_endpoint.scheme, parsed_endpoint.netloc, path)
        print("Sending to next endpoint: " + next_endpoint)
        code, task = self._post({'TaskId': taskId, 'Timestamp': self._now(), 'Status': 'created',
                                 'BackendStatus': 'created', 'Endpoint': next_endpoint, 'Body': body,
                                 'PublishToGrid': True})
        if code != 200:
            print("status code: " + str(code))
            return json.loads(_NOT_FOUND.format(taskId))
        return task

    def GetTaskStatus(self, taskId):
        if self.write_behind:
            self.flush()  # a read must see the writes this client has already accepted
        code, status = self._get(taskId)
        if code != 200:
            return json.loads(_NOT_FOUND.format(taskId))
        self._remember(taskId, status)
        return status
