"""SURVEY 8(f) row 2, the Redis half: `RedisTaskStore` runs the cache-connector state machine against a
Redis-protocol store with the reference's own keys (ProcessManager/CacheManager/CacheConnectorUpsert.cs:123-163,
CacheConnectorGet.cs:56, Libraries/QueueLogger.cs:44).  The server is tests/fake_redis.py (RESP2 over a real socket;
the image has no redis-server).  Checked: the wire client (binary-safe strings, nil, integers, arrays, errors,
reconnect), the keys and the single MULTI/EXEC round trip per upsert, the same scenario through `LocalTaskStore` and
`RedisTaskStore` giving the same answers, and failures (aborted transaction, store gone) mapping to 500."""
import json
import threading

import pytest

from task_management.local_task_store import LocalTaskStore
from task_management.redis_task_store import RedisTaskStore, RespClient, RespError
from tests.fake_redis import FakeRedis

ENDPOINT = "http://10.0.0.7/v1/org/classify"
PATH = "/v1/org/classify"


@pytest.fixture()
def redis():
    with FakeRedis() as r:
        yield r


def test_resp_client_round_trips(redis):
    c = RespClient(redis.host, redis.port)
    assert c.command("PING") == "PONG"
    blob = bytes(range(256)) + b"\r\n$5\r\n" + "ünï".encode()
    assert c.command("SET", "k", blob) == "OK"
    assert c.command("GET", "k") == blob                       # binary-safe, CRLF inside the value
    assert c.command("GET", "missing") is None                 # nil bulk
    assert c.command("ZADD", "z", 3, "c", 1, "a", 2, "b") == 3
    assert c.command("ZADD", "z", 5, "a") == 0                 # updated, not added
    assert c.command("ZRANGE", "z", 0, -1) == [b"b", b"c", b"a"]
    assert c.command("ZRANGE", "z", 0, -1, "WITHSCORES") == [b"b", b"2", b"c", b"3", b"a", b"5"]
    assert c.command("ZCARD", "z") == 3 and c.command("ZREM", "z", "b", "nope") == 1
    with pytest.raises(RespError, match="WRONGTYPE"):
        c.command("GET", "z")
    with pytest.raises(RespError, match="unknown command"):
        c.command("NOSUCH")
    out = c.pipeline([("SET", "a", "1"), ("GET", "a"), ("NOSUCH",), ("GET", "a")])
    assert out[0] == "OK" and out[1] == b"1" and isinstance(out[2], RespError) and out[3] == b"1"
    assert c.transaction([("SET", "t", "x"), ("ZADD", "zz", 1, "m")]) == ["OK", 1]
    redis.state.fail_exec = 1
    with pytest.raises(RespError, match="Unable to complete redis transaction"):
        c.transaction([("SET", "t", "y")])
    assert c.command("GET", "t") == b"x"
    # a dropped connection is re-opened once
    c._local.conn[0].close()
    assert c.command("PING") == "PONG"
    # one connection per thread
    seen = []
    t = threading.Thread(target=lambda: seen.append(c.command("GET", "t")))
    t.start()
    t.join()
    assert seen == [b"x"]


def test_upsert_writes_the_reference_keys_in_one_transaction(redis):
    store = RedisTaskStore(redis.host, redis.port)
    task = store.create_task(ENDPOINT, body="payload", publish=True)
    tid = task["TaskId"]
    names = [c[0] for c in redis.state.commands]
    assert names == [b"MULTI", b"SET", b"ZADD", b"SET", b"EXEC"] and store.round_trips == 1
    raw = RespClient(redis.host, redis.port)
    stored = json.loads(raw.command("GET", tid))
    assert set(stored) == {"TaskId", "Timestamp", "Status", "BackendStatus", "Endpoint", "Body", "PublishToGrid", "EndpointPath"}  # APITask.cs:12-27
    assert stored["Body"] is None and stored["EndpointPath"] == PATH and stored["BackendStatus"] == "created"
    assert raw.command("GET", tid + "_ORIG") == b"payload"                       # CacheConnectorUpsert.cs:38,158
    assert raw.command("ZSCORE", PATH + "_created", tid) is not None
    assert store.queue_length(PATH + "_created") == 1                            # what QueueLogger.cs:44 reads
    # running: out of created, into running - the same transaction
    del redis.state.commands[:]
    code, _ = store.upsert_task({"TaskId": tid, "Status": "running", "BackendStatus": "running", "Endpoint": ENDPOINT})
    assert code == 200 and [c[0] for c in redis.state.commands] == [b"MULTI", b"SET", b"ZADD", b"ZREM", b"EXEC"]
    assert redis.state.commands[3][1:] == [(PATH + "_created").encode(), tid.encode()]
    assert store.queue_length(PATH + "_created") == 0 and store.queue_length(PATH + "_running") == 1
    code, _ = store.upsert_task({"TaskId": tid, "Status": "completed", "BackendStatus": "completed", "Endpoint": ENDPOINT})
    assert code == 200 and store.set_members(PATH + "_running") == {} and tid in store.set_members(PATH + "_completed")
    assert store.get(tid)[0] == 200 and store.get("unknown") == (204, "")


def _scenario(store):
    """the same calls against either store -> what a caller can observe (ids and timestamps normalised away)"""
    published = []
    store.publisher = lambda task, body: published.append((task["BackendStatus"], body)) or (body != "refuse me")
    out = [store.upsert(""), store.upsert("   "), store.upsert("{not json"), store.upsert("[]"),
           store.upsert(json.dumps({"TaskId": "", "Status": "created"}))]                      # no Endpoint: upstream throws -> 500
    t = store.create_task(ENDPOINT, body="hello", publish=True)
    tid = t["TaskId"]
    code, again = store.upsert_task({"TaskId": tid, "Status": "created", "BackendStatus": "created", "Endpoint": ENDPOINT,
                                     "Body": "", "PublishToGrid": True})                     # a later publish re-reads the body
    out.append((code, again["TaskId"] == tid))
    code, text = store.upsert(json.dumps([{"TaskId": tid, "Status": "running x", "BackendStatus": "running", "Endpoint": ENDPOINT}]))
    out.append((code, {k: v for k, v in json.loads(text).items() if k != "Timestamp"}))
    out.append(sorted(store.set_members(PATH + "_running")) == [tid])
    out.append(store.set_members(PATH + "_created") == {})
    bad = store.create_task(ENDPOINT, body="refuse me", publish=True)                          # the publisher refuses: failed
    got = store.get_task(bad["TaskId"])
    out.append((got["BackendStatus"], got["Status"], bad["TaskId"] in store.set_members(PATH + "_failed"),
                bad["TaskId"] in store.set_members(PATH + "_created")))
    out.append(store.get("missing"))
    out.append([(s, b) for s, b in published])
    out = json.loads(json.dumps(out).replace(tid, "<tid>"))
    return out


def test_redis_store_and_local_store_answer_alike(redis):
    a = _scenario(LocalTaskStore())
    b = _scenario(RedisTaskStore(redis.host, redis.port))
    assert a == b
    assert a[0] == [400, ""] and a[2] == [500, ""] and a[4] == [500, ""]
    assert a[-1] == [["created", "hello"], ["created", "hello"], ["created", "refuse me"]]


def test_store_failures_are_500_not_exceptions(redis):
    store = RedisTaskStore(redis.host, redis.port)
    tid = store.create_task(ENDPOINT)["TaskId"]
    redis.state.fail_exec = 1
    assert store.upsert_task({"TaskId": tid, "Status": "running", "BackendStatus": "running", "Endpoint": ENDPOINT}) == (500, None)
    assert store.get_task(tid)["BackendStatus"] == "created"      # nothing was written
    redis.close()
    store.client.close()
    assert store.upsert_task({"TaskId": tid, "Status": "running", "BackendStatus": "running", "Endpoint": ENDPOINT}) == (500, None)
    assert store.get(tid) == (500, "")


def test_many_threads_keep_the_sets_consistent(redis):
    store = RedisTaskStore(redis.host, redis.port)
    ids = [store.create_task(ENDPOINT)["TaskId"] for _ in range(40)]

    def work(part):
        for tid in part:
            for status in ("running", "completed"):
                assert store.upsert_task({"TaskId": tid, "Status": status, "BackendStatus": status, "Endpoint": ENDPOINT})[0] == 200

    threads = [threading.Thread(target=work, args=(ids[k::4],)) for k in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert store.queue_length(PATH + "_created") == 0 and store.queue_length(PATH + "_running") == 0
    assert sorted(store.set_members(PATH + "_completed")) == sorted(ids)
