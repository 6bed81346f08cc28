"""In-process task store that honours the contract of the platform's cache-connector
functions, so the hot path can be driven and benchmarked "in-process without network"
(BASELINE.json north_star).  The Azure control plane itself is out of scope.

Contract restated from the reference (citations relative to /root/reference):
  * task JSON fields TaskId, Timestamp, Status, BackendStatus, Endpoint, Body, PublishToGrid
    (ProcessManager/Classes/APITask.cs:12-18);
  * upsert: an empty body is a 400 (CacheManager/CacheConnectorUpsert.cs:57-74), a body that does not
    deserialise a 500 (:76-82), an array body means its first element (:60-63); an empty
    TaskId gets a new GUID (:92-100); Timestamp is overwritten with UTC now (:102); the stored
    JSON has Body stripped (:119-127); the task id is added to the sorted set
    "{EndpointPath}_{BackendStatus}" scored by epoch seconds (:130-133) and removed from the
    previous state's set - running removes it from created, completed/failed from running
    (:135-142); with PublishToGrid the original body is kept under "{TaskId}_ORIG" (:158) and a
    later publish without body re-reads it (:175); the response is the stored JSON;
  * get: 200 + JSON, or 204 when the id is unknown (CacheManager/CacheConnectorGet.cs:56-66).
"""
from __future__ import annotations

import datetime
import json
import threading
import time
import uuid
from urllib.parse import urlparse

BACKEND_STATUSES = ("created", "running", "completed", "failed")
ORIG_SUFFIX = "_ORIG"


_stamp_cache = (0, "")


def _utc_stamp(now: int) -> str:
    """UTC "MM/DD/YYYY HH:MM:SS" of an epoch second; formatted once per second (strftime is a tenth of an upsert)"""
    global _stamp_cache
    sec, text = _stamp_cache
    if sec != now:
        text = datetime.datetime.fromtimestamp(now, datetime.timezone.utc).strftime("%m/%d/%Y %H:%M:%S")
        _stamp_cache = (now, text)
    return text


def prepare_upsert(task, path_cache: dict):
    """what CacheConnectorUpsert does to the request before it touches the store (:92-127): new GUID for an empty
    TaskId, Timestamp = UTC now, Body stripped from what is stored, EndpointPath from the Endpoint URI.
    -> (task to store, the request's body, epoch seconds for the sorted-set score), or None for a task the
    upstream function would throw on (no Endpoint / BackendStatus: its catch answers 500)."""
    if not isinstance(task, dict) or not task.get("Endpoint") or not task.get("BackendStatus"):
        return None  # upstream: Uri(null) / BackendStatus.Equals on null throw inside the try (:130-142, :204-210)
    task = dict(task)
    task_body = task.get("Body")
    if not task.get("TaskId"):
        task["TaskId"] = str(uuid.uuid4())
    now = int(time.time())
    task["Timestamp"] = _utc_stamp(now)
    task["Body"] = None
    task.setdefault("Status", None)
    task.setdefault("PublishToGrid", False)
    endpoint = task["Endpoint"]
    path = path_cache.get(endpoint)
    if path is None:
        path = path_cache[endpoint] = urlparse(endpoint).path
    task["EndpointPath"] = path  # a public getter of APITask, so the serialiser emits it (Classes/APITask.cs:20-27)
    return task, task_body, now


class LocalTaskStore:
    def __init__(self, publisher=None):
        self._lock = threading.Lock()
        self.strings: dict[str, str] = {}        # "{TaskId}_ORIG" -> original body
        self.tasks: dict[str, dict] = {}         # TaskId -> stored task (what Redis holds as JSON text)
        self._paths: dict[str, str] = {}         # Endpoint -> its path (urlparse is the costly part of an upsert)
        self.sorted_sets: dict[str, dict[str, int]] = {}
        self.publisher = publisher  # callable(task_dict, body) -> bool, the "Event Grid / Service Bus" stand-in
        self.upserts = 0
        self.gets = 0

    # ---- the two HTTP functions, as (status_code, text)
    def upsert(self, body_text: str | None):
        """the HTTP function: JSON text in, (status, JSON text) out."""
        if not body_text or not body_text.strip():
            self.upserts += 1
            return 400, ""
        try:
            task = json.loads(body_text)
            if isinstance(task, list):  # upstream takes the first element of an array body (:60-63)
                task = task[0]
        except (ValueError, IndexError):
            self.upserts += 1
            return 500, ""  # upstream: the deserialiser throws and the function answers 500 (:76-82)
        code, stored = self.upsert_task(task)
        return code, (json.dumps(stored) if code == 200 else "")

    def upsert_task(self, task):
        """the same state machine for a caller in this process that already holds the task as a dict (no
        JSON round trip): -> (status, stored task dict)."""
        self.upserts += 1
        prepared = prepare_upsert(task, self._paths)
        if prepared is None:
            return 500, None
        task, task_body, now = prepared
        path, status, tid = task["EndpointPath"], task["BackendStatus"], task["TaskId"]
        publish = bool(task.get("PublishToGrid"))
        with self._lock:
            self.tasks[tid] = task
            self.sorted_sets.setdefault(f"{path}_{status}", {})[tid] = now
            if status == "running":
                self.sorted_sets.get(f"{path}_created", {}).pop(tid, None)
            elif status in ("completed", "failed"):
                self.sorted_sets.get(f"{path}_running", {}).pop(tid, None)
            if publish:
                if task_body:
                    self.strings[tid + ORIG_SUFFIX] = task_body if isinstance(task_body, str) else json.dumps(task_body)
                else:
                    task_body = self.strings.get(tid + ORIG_SUFFIX)
        if publish and self.publisher is not None:
            ok = False
            try:
                ok = bool(self.publisher(dict(task), task_body))
            except Exception:  # noqa: BLE001 - a failed publish moves the task to failed, as upstream
                ok = False
            if not ok:
                failed = dict(task)
                failed["Status"] = "Failed - unable to send to backend service."
                failed["BackendStatus"] = "failed"
                with self._lock:
                    self.tasks[tid] = failed
                    self.sorted_sets.setdefault(f"{path}_failed", {})[tid] = now
                    self.sorted_sets.get(f"{path}_created", {}).pop(tid, None)
        return 200, task  # (upstream answers with what it serialised before publishing, also when the publish failed)

    def get(self, task_id: str):
        """the HTTP function: (200, JSON text) or (204, '')."""
        t = self.get_task(task_id)
        return (200, json.dumps(t)) if t is not None else (204, "")

    def get_task(self, task_id: str):
        """-> a copy of the stored task dict, or None (no JSON round trip)."""
        self.gets += 1
        with self._lock:
            t = self.tasks.get(task_id)
        return dict(t) if t is not None else None

    # ---- helpers for tests / the in-process platform
    def create_task(self, endpoint: str, body: str = "", publish: bool = False) -> dict:
        """what APIM's request policy does (APIManagement/request_policy.xml:5-21)."""
        code, task = self.upsert_task({"TaskId": "", "Status": "created", "BackendStatus": "created",
                                       "Endpoint": endpoint, "Body": body, "PublishToGrid": publish})
        assert code == 200
        return dict(task)

    def set_members(self, key: str):
        with self._lock:
            return dict(self.sorted_sets.get(key, {}))


_default_store = None
_default_lock = threading.Lock()


def default_store() -> LocalTaskStore:
    global _default_store
    with _default_lock:
        if _default_store is None:
            _default_store = LocalTaskStore()
        return _default_store
