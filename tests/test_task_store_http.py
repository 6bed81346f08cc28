"""the task store over its real wire contract.  An HTTP facade of the two
cache-connector functions (state machine of ProcessManager/CacheManager/CacheConnectorUpsert.cs:92-212
and CacheConnectorGet.cs:53-72, JSON of ProcessManager/Classes/APITask.cs:12-28) is driven by

  * the reference's OWN client, unmodified (APIs/1.0/Common/task_management/distributed_api_task.py,
    aiohttp), when /root/reference is present - the facade must satisfy it;
  * this repo's client over its `requests` transport, plain, coalescing and write-behind.
"""
import importlib.util
import json
import os

import pytest

from task_management.cache_connector_http import serve
from task_management.local_task_store import LocalTaskStore

REF_CLIENT = "/root/reference/APIs/1.0/Common/task_management/distributed_api_task.py"
ENDPOINT = "http://10.0.0.7/v1/org/classify"


@pytest.fixture(params=["memory", "redis"])
def served(monkeypatch, request):
    """the facade over the in-process store and over the Redis-protocol store (tests/fake_redis.py on a socket)"""
    import contextlib
    with contextlib.ExitStack() as stack:
        if request.param == "redis":
            from task_management.redis_task_store import RedisTaskStore
            from tests.fake_redis import FakeRedis
            r = stack.enter_context(FakeRedis())
            store = RedisTaskStore(r.host, r.port)
        else:
            store = LocalTaskStore()
        srv = stack.enter_context(serve(store))
        monkeypatch.setenv("CACHE_CONNECTOR_UPSERT_URI", srv.upsert_uri)
        monkeypatch.setenv("CACHE_CONNECTOR_GET_URI", srv.get_uri)
        monkeypatch.delenv("AI4E_TASK_STORE", raising=False)
        monkeypatch.delenv("AI4E_TASK_COALESCE", raising=False)
        monkeypatch.delenv("AI4E_TASK_WRITE_BEHIND", raising=False)
        yield srv


def _lifecycle(client, store, flush=lambda: None):
    """created (platform) -> running -> completed, then a pipeline hop; returns the task id."""
    assert client.AddTask() == {"TaskId": "-1", "Status": "error"}       # containers do not create tasks (400)
    assert client.GetTaskStatus("nope") == {"TaskId": "nope", "Status": "not found"}   # 204
    tid = store.create_task(ENDPOINT, body="payload", publish=True)["TaskId"]
    got = client.GetTaskStatus(tid)
    assert got["TaskId"] == tid and got["BackendStatus"] == "created" and got["Endpoint"] == ENDPOINT
    assert got["Body"] is None and got["EndpointPath"] == "/v1/org/classify"
    r = client.UpdateTaskStatus(tid, "running model")
    assert r["TaskId"] == tid and r["Status"] == "running model" and r["BackendStatus"] == "running"
    flush()
    assert tid in store.set_members("/v1/org/classify_running") and tid not in store.set_members("/v1/org/classify_created")
    r = client.CompleteTask(tid, "completed - [1, 2, 3]")
    assert r["BackendStatus"] == "completed" and r["Endpoint"] == ENDPOINT
    flush()
    assert tid in store.set_members("/v1/org/classify_completed") and tid not in store.set_members("/v1/org/classify_running")
    stored = json.loads(store.get(tid)[1])
    assert stored["Status"] == "completed - [1, 2, 3]" and stored["PublishToGrid"] is False
    return tid


@pytest.mark.skipif(not os.path.exists(REF_CLIENT), reason="the reference tree is not on this machine")
def test_unmodified_reference_client_round_trips(served):
    spec = importlib.util.spec_from_file_location("ref_distributed_api_task", REF_CLIENT)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    client = mod.DistributedApiTaskManager()
    published = []
    served.store.publisher = lambda task, body: published.append((task["Endpoint"], body)) or True
    tid = _lifecycle(client, served.store)
    # the reference pays a read before every write: AddTask-less lifecycle = 2 x (GET + POST) + 2 reads above
    assert served.facade.requests["get"] >= 4 and served.facade.requests["upsert"] >= 3
    nxt = client.AddPipelineTask(tid, "org", "v2", "detect", "next-body")
    assert nxt["Endpoint"] == "http://10.0.0.7/v2/org/detect" and nxt["BackendStatus"] == "created"
    assert published[-1] == ("http://10.0.0.7/v2/org/detect", "next-body")
    assert tid in served.store.set_members("/v2/org/detect_created")
    client.FailTask(tid, "Task failed - try again")
    assert json.loads(served.store.get(tid)[1])["BackendStatus"] == "failed"


def test_this_repos_client_over_requests(served):
    from task_management.distributed_api_task import DistributedApiTaskManager
    client = DistributedApiTaskManager()
    assert client.store is None            # the HTTP transport, not the in-process store
    g0, u0 = served.facade.requests["get"], served.facade.requests["upsert"]
    tid = _lifecycle(client, served.store)
    # same traffic as the reference: every status write is a read + a write
    assert served.facade.requests["get"] - g0 == 2 + 2 and served.facade.requests["upsert"] - u0 == 1 + 2
    nxt = client.AddPipelineTask(tid, "org", "v2", "detect", "next-body")
    assert nxt["Endpoint"] == "http://10.0.0.7/v2/org/detect"
    # a dead platform never raises into the task function: sentinel dicts, like upstream
    served.close()
    assert client.GetTaskStatus(tid) == {"TaskId": tid, "Status": "not found"}
    assert client.UpdateTaskStatus(tid, "x") == {"TaskId": tid, "Status": "not found"}


def test_coalescing_client_writes_without_reading(served):
    from task_management.distributed_api_task import DistributedApiTaskManager
    client = DistributedApiTaskManager(coalesce=True)
    tid = served.store.create_task(ENDPOINT)["TaskId"]
    assert client.GetTaskStatus(tid)["Endpoint"] == ENDPOINT       # what TaskManager.AddTask does for a taskId header
    g0, u0 = served.facade.requests["get"], served.facade.requests["upsert"]
    client.UpdateTaskStatus(tid, "running")
    client.CompleteTask(tid, "completed - ok")
    assert served.facade.requests["get"] == g0                       # no read-before-write
    assert served.facade.requests["upsert"] - u0 == 2                # one upsert per status change
    stored = json.loads(served.store.get(tid)[1])
    assert stored["BackendStatus"] == "completed" and stored["Endpoint"] == ENDPOINT
    # an id this client never read still takes the reference's read-modify-write path
    other = served.store.create_task(ENDPOINT)["TaskId"]
    client.UpdateTaskStatus(other, "running")
    assert served.facade.requests["get"] == g0 + 1


def test_write_behind_keeps_order_and_flushes(served):
    from task_management.distributed_api_task import DistributedApiTaskManager
    client = DistributedApiTaskManager(write_behind=True)
    tids = [served.store.create_task(ENDPOINT)["TaskId"] for _ in range(20)]
    for t in tids:
        client.GetTaskStatus(t)
    for t in tids:
        r = client.UpdateTaskStatus(t, "running")
        assert r["BackendStatus"] == "running" and r["TaskId"] == t      # answered at once, from the payload
        client.CompleteTask(t, "completed - " + t)
    client.flush()
    for t in tids:
        stored = json.loads(served.store.get(t)[1])
        assert stored["BackendStatus"] == "completed" and stored["Status"] == "completed - " + t
    assert len(served.store.set_members("/v1/org/classify_completed")) == 20
    assert not served.store.set_members("/v1/org/classify_running")


def test_facade_status_codes(served):
    import requests
    assert requests.post(served.upsert_uri, data="").status_code == 400
    assert requests.post(served.upsert_uri, data="{not json").status_code == 500
    r = requests.post(served.upsert_uri, data=json.dumps([{"TaskId": "", "Status": "created", "BackendStatus": "created",
                                                           "Endpoint": ENDPOINT, "Body": "b", "PublishToGrid": False}]))
    assert r.status_code == 200 and len(r.json()["TaskId"]) == 36        # an array body: first element; new GUID
    assert requests.get(served.get_uri, params={"taskId": r.json()["TaskId"]}).status_code == 200
    assert requests.get(served.get_uri, params={"taskId": "missing"}).status_code == 204
    assert requests.get(served.get_uri).status_code == 204               # no taskId: looks up "nil"
    assert requests.get(served.base + "/api/Other").status_code == 404
