"""CPU tests of the drop-in Python surface (boundary B1).

`test_matches_reference_trace` compares this repo's `ai4e_service.APIService` with what the
UNMODIFIED reference class did in the same scripted conversation (fixture
tests/golden/host_shell_trace.json, minted by tests/golden/make_host_shell_trace.py);
`test_reference_still_produces_the_trace` re-runs the reference where /root/reference exists."""
import json
import os
import subprocess
import sys
import threading
import time

import pytest

from tests.host_shell_scenario import RecordingTaskManager, run_scenario

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TRACE = os.path.join(ROOT, "tests", "golden", "host_shell_trace.json")


@pytest.fixture(autouse=True)
def _clean_env(monkeypatch):
    for k in ("API_PREFIX", "AI4E_BUSY_STATUS", "CACHE_CONNECTOR_UPSERT_URI", "CACHE_CONNECTOR_GET_URI"):
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setenv("AI4E_TASK_STORE", "local")


def _flask():
    try:
        from flask import Flask
    except ImportError:
        from ai4e_b200.compat.miniflask import Flask
    return Flask


def test_matches_reference_trace():
    from ai4e_service import APIService
    want = json.load(open(TRACE))
    want.pop("_source")
    got = run_scenario(APIService, _flask())
    for key in sorted(want):
        assert got[key] == want[key], f"{key}: ours {got[key]!r} != reference {want[key]!r}"


@pytest.mark.skipif(not os.path.exists("/root/reference/APIs/1.0/base-py/ai4e_service.py"),
                    reason="the reference tree only exists in the build container")
def test_reference_still_produces_the_trace():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "golden", "make_host_shell_trace.py"), "--print"],
                       capture_output=True, text=True, cwd=ROOT, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    live = json.loads(r.stdout.strip().splitlines()[-1])
    assert live == json.load(open(TRACE))


def test_prefix_admission_and_busy_status(monkeypatch):
    """documented differences: checks fire under API_PREFIX; AI4E_BUSY_STATUS=429 is opt-in."""
    monkeypatch.setenv("API_PREFIX", "/v1/org")
    monkeypatch.setenv("AI4E_BUSY_STATUS", "429")
    from ai4e_service import APIService
    app = _flask()("p")
    svc = APIService(app, type("L", (), {"tracer": None, "log_exception": lambda *a, **k: None})())

    @svc.api_sync_func(api_path="/full", methods=["GET"], maximum_concurrent_requests=0)
    def full(*a, **k):
        return "x"

    @svc.api_sync_func(api_path="/open", methods=["POST"], content_types=["image/jpeg"])
    def open_(*a, **k):  # maximum_concurrent_requests=None: no limit
        return "ok"

    c = app.test_client()
    assert c.get("/v1/org/").get_data(as_text=True) == "Health check OK"
    assert c.get("/v1/org/full").status_code == 429
    assert c.post("/v1/org/open", data=b"1", headers={"Content-Type": "text/plain"}).status_code == 401
    assert c.post("/v1/org/open", data=b"1", headers={"Content-Type": "image/jpeg"}).get_data(as_text=True) == "ok"


def test_tracer_span_wraps_the_endpoint():
    from ai4e_service import APIService
    spans = []

    class Tracer:
        def span(self, name):
            import contextlib

            @contextlib.contextmanager
            def cm():
                spans.append(("enter", name))
                yield object()
                spans.append(("exit", name))
            return cm()

    app = _flask()("t")
    svc = APIService(app, type("L", (), {"tracer": Tracer(), "log_exception": lambda *a, **k: None})())

    @svc.api_sync_func(api_path="/a", methods=["GET"], maximum_concurrent_requests=2, trace_name="get:a")
    def a(*args, **kwargs):
        spans.append(("func", None))
        return "a"

    @svc.api_sync_func(api_path="/b", methods=["GET"], maximum_concurrent_requests=2)
    def b(*args, **kwargs):
        return "b"

    c = app.test_client()
    c.get("/a"); c.get("/b")
    assert spans == [("enter", "get:a"), ("func", None), ("exit", "get:a"), ("enter", "/b"), ("exit", "/b")]


# ------------------------------------------------------------------ task store contract
def test_local_store_state_machine():
    from task_management.local_task_store import LocalTaskStore
    published = []
    st = LocalTaskStore(publisher=lambda task, body: published.append((task["TaskId"], body)) or True)
    assert st.upsert("")[0] == 400 and st.upsert(None)[0] == 400          # empty body -> 400
    assert st.get("nope") == (204, "")                                     # unknown id -> 204
    t = st.create_task("http://10.0.0.1/v1/org/classify", body="payload", publish=True)
    tid = t["TaskId"]
    assert len(tid) == 36 and t["Body"] is None                            # new GUID, body stripped
    assert published == [(tid, "payload")]
    assert st.strings[tid + "_ORIG"] == "payload"
    assert tid in st.set_members("/v1/org/classify_created")
    code, txt = st.upsert(json.dumps({"TaskId": tid, "Status": "running - 1/2", "BackendStatus": "running",
                                      "Endpoint": t["Endpoint"], "PublishToGrid": False}))
    assert code == 200 and json.loads(txt)["Status"] == "running - 1/2"
    assert tid in st.set_members("/v1/org/classify_running") and tid not in st.set_members("/v1/org/classify_created")
    st.upsert(json.dumps({"TaskId": tid, "Status": "completed - [1,2]", "BackendStatus": "completed",
                          "Endpoint": t["Endpoint"], "PublishToGrid": False}))
    assert tid in st.set_members("/v1/org/classify_completed") and tid not in st.set_members("/v1/org/classify_running")
    assert json.loads(st.get(tid)[1])["BackendStatus"] == "completed"
    # a pipeline republish without body re-reads the original body
    st.upsert(json.dumps({"TaskId": tid, "Status": "created", "BackendStatus": "created",
                          "Endpoint": "http://10.0.0.1/v1/org/next", "PublishToGrid": True}))
    assert published[-1] == (tid, "payload")


def test_local_store_failed_publish_moves_task_to_failed():
    from task_management.local_task_store import LocalTaskStore
    st = LocalTaskStore(publisher=lambda task, body: False)
    t = st.create_task("http://h/v1/a/b", body="x", publish=True)
    stored = json.loads(st.get(t["TaskId"])[1])
    assert stored["BackendStatus"] == "failed" and stored["Status"].startswith("Failed - unable to send")
    assert t["TaskId"] in st.set_members("/v1/a/b_failed") and t["TaskId"] not in st.set_members("/v1/a/b_created")


def test_task_manager_contract_over_local_store():
    from task_management.api_task import TaskManager
    from task_management.local_task_store import LocalTaskStore
    st = LocalTaskStore()
    tm = TaskManager(store=st)
    # no header: the distributed build cannot create tasks (platform answers 400) -> sentinel
    assert tm.AddTask(None) == {"TaskId": "-1", "Status": "error"}
    t = st.create_task("http://h/v1/org/classify")
    req = type("R", (), {"headers": {"taskId": t["TaskId"]}})()
    assert tm.AddTask(req)["TaskId"] == t["TaskId"]
    assert tm.GetTaskStatus("missing") == {"TaskId": "missing", "Status": "not found"}
    r = tm.UpdateTaskStatus(t["TaskId"], "running model")
    assert r["BackendStatus"] == "running" and r["Endpoint"] == "http://h/v1/org/classify" and r["PublishToGrid"] is False
    assert tm.CompleteTask(t["TaskId"], "completed - ok")["BackendStatus"] == "completed"
    assert tm.FailTask(t["TaskId"], "Task failed - try again") is None
    assert tm.GetTaskStatus(t["TaskId"])["BackendStatus"] == "failed"
    # every status write is a read-modify-write like upstream (1 GET + 1 POST)
    g0, u0 = st.gets, st.upserts
    tm.UpdateTaskStatus(t["TaskId"], "again")
    assert (st.gets - g0, st.upserts - u0) == (1, 1)
    # a status write for an unknown id is created against http://localhost, as upstream
    assert tm.UpdateTaskStatus("ghost", "running")["Endpoint"] == "http://localhost"
    # pipelining rewrites the endpoint path and republishes
    nxt = tm.AddPipelineTask(t["TaskId"], "org", "v2", "segment", "body2")
    assert nxt["Endpoint"] == "http://h/v2/org/segment" and nxt["BackendStatus"] == "created"
    assert tm.AddPipelineTask("missing", "org", "v2", "segment", "b") == {"TaskId": "-1", "Status": "error"}


def test_timestamp_format_of_status_writes():
    import re
    from task_management.distributed_api_task import DistributedApiTaskManager
    assert re.fullmatch(r"\d{4}-\d\d-\d\d \d\d:\d\d:\d\d", DistributedApiTaskManager._now())


def test_in_process_platform_round_trip():
    """APIM creates the task -> the store publishes -> the 'queue pusher' POSTs to the container with
    the taskId header -> the async endpoint runs the task function on a worker -> status lands in
    the store.  No network, no Azure: BASELINE.json benchmarks the control plane in-process."""
    from ai4e_service import APIService
    from task_management.api_task import TaskManager
    from task_management.local_task_store import LocalTaskStore
    app = _flask()("platform")
    client = app.test_client()
    pushed = []

    def pusher(task, body):  # BackendQueueProcessor.cs:48-52 in one line
        r = client.post("/v1/org/classify", data=body.encode(), headers={"taskId": task["TaskId"],
                                                                         "Content-Type": "application/octet-stream"})
        pushed.append(r.status_code)
        return r.status_code == 200

    store = LocalTaskStore(publisher=pusher)
    svc = APIService(app, type("L", (), {"tracer": None, "log_exception": lambda *a, **k: None})())
    svc.api_task_manager = TaskManager(store=store)
    done = threading.Event()

    @svc.api_async_func(api_path="/v1/org/classify", methods=["POST"], maximum_concurrent_requests=64,
                        request_processing_function=lambda request: {"payload": request.get_data()})
    def classify(*args, **kwargs):
        tid = kwargs["taskId"]
        svc.api_task_manager.UpdateTaskStatus(tid, "running")
        assert kwargs["request"].headers["taskId"] == tid and kwargs["request"].get_data() == kwargs["payload"]
        svc.api_task_manager.CompleteTask(tid, "completed - " + json.dumps({"top5": [1, 2, 3, 4, 5], "n": len(kwargs["payload"])}))
        done.set()

    task = store.create_task("http://container/v1/org/classify", body="raw-image-bytes", publish=True)
    assert done.wait(5) and pushed == [200]
    for _ in range(100):
        final = json.loads(store.get(task["TaskId"])[1])
        if final["BackendStatus"] == "completed":
            break
        time.sleep(0.01)
    assert final["BackendStatus"] == "completed" and json.loads(final["Status"][len("completed - "):])["n"] == 15
    assert task["TaskId"] in store.set_members("/v1/org/classify_completed")
    assert client.get("/task/" + task["TaskId"]).get_json()["Status"].startswith("completed - ")


def test_worker_pool_runs_many_tasks_concurrently():
    """64 tasks blocked at the same time: what lets the engine's micro-batcher fill a batch."""
    from ai4e_service import APIService
    app = _flask()("w")
    svc = APIService(app, type("L", (), {"tracer": None, "log_exception": lambda *a, **k: None})())
    tm = RecordingTaskManager()
    svc.api_task_manager = tm
    n = 64
    gate, inside = threading.Event(), threading.Semaphore(0)

    @svc.api_async_func(api_path="/c", methods=["POST"], maximum_concurrent_requests=n)
    def c_(*args, **kwargs):
        inside.release()
        gate.wait(10)

    cl = app.test_client()
    for i in range(n):
        assert cl.post("/c", headers={"taskId": str(i)}).get_data(as_text=True) == "TaskId: %d" % i
    for _ in range(n):
        assert inside.acquire(timeout=10)
    assert svc.func_request_counts["/c"] == n
    assert cl.post("/c", headers={"taskId": "x"}).status_code == 503  # admission: over the limit
    gate.set()
    for _ in range(500):
        if svc.func_request_counts["/c"] == 0:
            break
        time.sleep(0.01)
    assert svc.func_request_counts["/c"] == 0


def test_back_to_back_tasks_do_not_share_one_idle_worker():
    """Two async requests that arrive while one worker is idle must both start at once (the reference
    starts a thread per task, ai4e_service.py:180-183): the second may not wait for the first."""
    from ai4e_service import _TaskWorkers
    w = _TaskWorkers(16)
    warm = threading.Event()
    w.run(warm.set, (), {})
    assert warm.wait(5)
    time.sleep(0.05)  # the worker is idle again
    started, gate = [], threading.Event()

    def job(name):
        started.append((name, time.perf_counter()))
        gate.wait(5)

    t0 = time.perf_counter()
    w.run(job, ("a",), {})
    w.run(job, ("b",), {})
    for _ in range(200):
        if len(started) == 2:
            break
        time.sleep(0.005)
    gate.set()
    assert len(started) == 2 and max(t for _, t in started) - t0 < 0.5


class _FakePending:
    """the surface of ai4e_b200.engine.Pending that the shell relies on."""

    def __init__(self):
        self.error, self._cbs, self.done = None, [], False

    def add_done_callback(self, fn):
        self._cbs.append(fn)

    def resolve(self, error=None):
        self.error = error
        for fn in self._cbs:
            fn(self)


def test_deferred_task_keeps_the_request_counted_and_fails_like_an_exception():
    """A task function that returns a Pending (engine.classify_later) frees its worker at once; the
    request stays in flight until the Pending resolves, and an error there becomes
    FailTask(taskId, 'Task failed - try again') like any exception (reference :185-195)."""
    from ai4e_service import APIService
    app = _flask()("d")
    svc = APIService(app, type("L", (), {"tracer": None, "log_exception": lambda *a, **k: None})())
    tm = RecordingTaskManager()
    svc.api_task_manager = tm
    pendings = {}

    @svc.api_async_func(api_path="/c", methods=["POST"], maximum_concurrent_requests=8)
    def c_(*args, **kwargs):
        p = pendings[kwargs["taskId"]] = _FakePending()
        return p

    cl = app.test_client()
    for tid in ("ok", "bad"):
        assert cl.post("/c", headers={"taskId": tid}).get_data(as_text=True) == "TaskId: " + tid
    for _ in range(200):
        if len(pendings) == 2:
            break
        time.sleep(0.005)
    time.sleep(0.05)
    assert svc.func_request_counts["/c"] == 2          # the functions returned, the tasks have not finished
    pendings["ok"].resolve()
    assert svc.func_request_counts["/c"] == 1
    pendings["bad"].resolve(RuntimeError("GPU batch failed"))
    assert svc.func_request_counts["/c"] == 0
    fails = [c for c in tm.calls if c[0] == "FailTask"]
    assert fails == [["FailTask", "bad", "Task failed - try again"]]


def test_inline_async_endpoint_runs_on_the_request_thread():
    """`api_async_func(..., inline=True)`: a non-blocking task function (one that returns a Pending) runs on the
    request thread - same response, same kwargs, same accounting and failure mapping, no worker hand-off."""
    from ai4e_service import APIService
    app = _flask()("i")
    svc = APIService(app, type("L", (), {"tracer": None, "log_exception": lambda *a, **k: None})())
    tm = RecordingTaskManager()
    svc.api_task_manager = tm
    seen, pend = {}, {}

    @svc.api_async_func(api_path="/c", methods=["POST"], maximum_concurrent_requests=4, inline=True,
                        request_processing_function=lambda request: {"n": len(request.get_data())})
    def c_(*args, **kwargs):
        seen[kwargs["taskId"]] = (threading.current_thread(), sorted(kwargs), kwargs["request"].headers.get("taskId"))
        if kwargs["taskId"] == "boom":
            raise RuntimeError("x")
        p = pend[kwargs["taskId"]] = _FakePending()
        return p

    cl = app.test_client()
    assert cl.post("/c", data=b"abc", headers={"taskId": "a"}).get_data(as_text=True) == "TaskId: a"
    assert seen["a"][0] is threading.current_thread()                 # ran before the response, on this thread
    assert seen["a"][1] == ["api_path", "func", "n", "request", "taskId"] and seen["a"][2] == "a"
    assert svc.func_request_counts["/c"] == 1                          # still in flight: the Pending is open
    pend["a"].resolve()
    assert svc.func_request_counts["/c"] == 0
    assert cl.post("/c", data=b"", headers={"taskId": "boom"}).get_data(as_text=True) == "TaskId: boom"
    assert [c for c in tm.calls if c[0] == "FailTask"] == [["FailTask", "boom", "Task failed - try again"]]
    assert svc.func_request_counts["/c"] == 0


def test_metrics_endpoint_uses_the_platform_metric_names(monkeypatch):
    """in-flight requests under the name the HPA consumes
    (CURRENT_REQUESTS/{cluster}{prefix}{path}), counted while a request is being served, plus the
    attached engine's queue depth and latency."""
    import json
    monkeypatch.setenv("API_PREFIX", "/v1/org")
    from ai4e_service import APIService
    app = _flask()("m")
    svc = APIService(app, type("L", (), {"tracer": None, "log_exception": lambda *a, **k: None})())
    seen = {}

    @svc.api_sync_func(api_path="/classify", methods=["POST"], maximum_concurrent_requests=4)
    def classify(*a, **k):
        seen.update(svc.metrics())  # taken while this request is in flight
        return "ok"

    class FakeEngine:
        def stats(self):
            return {"queue_depth": 7, "images": 100, "batches": 3, "lat_ms_p50": 1.5, "lat_ms_p99": 2.5, "healthy": 1}

    svc.attach_engine(FakeEngine())  # also adds the /metrics route: before the first request
    c = app.test_client()
    assert c.post("/v1/org/classify", data=b"x").get_data(as_text=True) == "ok"
    import ai4e_service
    name = "CURRENT_REQUESTS/%s/v1/org/classify" % ai4e_service.service_cluster
    assert seen[name] == 1
    m = json.loads(c.get("/v1/org/metrics").get_data(as_text=True))
    assert m[name] == 0
    assert m["ENGINE/%s/replica0/queue_depth" % ai4e_service.service_cluster] == 7
    assert m["ENGINE/%s/replica0/lat_ms_p99" % ai4e_service.service_cluster] == 2.5
