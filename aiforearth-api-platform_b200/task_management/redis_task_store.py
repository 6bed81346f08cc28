"""The cache-connector state machine against a Redis-protocol store, key for key what the
platform's C# functions read and write, so a task written here is a task `CacheConnectorGet`,
`BackendQueueProcessor` and `QueueLogger` of the reference would find:

    SET   {TaskId}                          the APITask JSON, Body stripped          (CacheConnectorUpsert.cs:119-127)
    ZADD  {EndpointPath}_{BackendStatus}    member = TaskId, score = epoch seconds   (:130-133)
    ZREM  {EndpointPath}_created            when the task turns running              (:135-138)
    ZREM  {EndpointPath}_running            when it completes or fails               (:139-142)
    SET   {TaskId}_ORIG                     the original body of a published task    (:38, :158), re-read at :175
  all in one MULTI / EXEC, as the upstream transaction (:123, :163);
    GET   {TaskId}                          CacheConnectorGet.cs:56, BackendQueueProcessor.cs:98
    ZCARD {EndpointPath}_{BackendStatus}    Libraries/QueueLogger.cs:44 (queue lengths for the autoscaler)
(citations relative to /root/reference/ProcessManager).  Same public methods as `LocalTaskStore`, so the HTTP
facade (`cache_connector_http.serve(RedisTaskStore(...))`), the drop-in task manager and the tests take either.

The client below speaks RESP2 over a plain socket (redis-py is not a dependency of the base image): one
connection per thread, commands pipelined per transaction - one network round trip per upsert.
"""
from __future__ import annotations

import json
import socket
import threading

from task_management.local_task_store import ORIG_SUFFIX, prepare_upsert


class RespError(Exception):
    """an error reply (`-ERR ...`) or a broken connection"""


class RespClient:
    """Minimal RESP2 client: `command(*args)`, `pipeline([...])`, `transaction([...])` (MULTI ... EXEC)."""

    def __init__(self, host: str = "127.0.0.1", port: int = 6379, timeout: float = 5.0, db: int = 0, password: str | None = None):
        self.host, self.port, self.timeout, self.db, self.password = host, port, timeout, db, password
        self._local = threading.local()

    # ---- wire
    @staticmethod
    def _encode(args) -> bytes:
        out = [b"*%d\r\n" % len(args)]
        for a in args:
            b = a if isinstance(a, (bytes, bytearray)) else str(a).encode()
            out.append(b"$%d\r\n" % len(b))
            out.append(b)
            out.append(b"\r\n")
        return b"".join(out)

    def _conn(self):
        c = getattr(self._local, "conn", None)
        if c is None:
            sock = socket.create_connection((self.host, self.port), timeout=self.timeout)
            sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            c = self._local.conn = (sock, sock.makefile("rb"))
            if self.password:
                self._roundtrip(c, [("AUTH", self.password)])
            if self.db:
                self._roundtrip(c, [("SELECT", self.db)])
        return c

    def close(self):
        c = getattr(self._local, "conn", None)
        if c is not None:
            self._local.conn = None
            try:
                c[1].close()
                c[0].close()
            except OSError:
                pass

    @classmethod
    def _read(cls, f):
        line = f.readline()
        if not line.endswith(b"\r\n"):
            raise RespError("connection closed by the store")
        kind, rest = line[:1], line[1:-2]
        if kind == b"+":
            return rest.decode()
        if kind == b"-":
            return RespError(rest.decode())
        if kind == b":":
            return int(rest)
        if kind == b"$":
            n = int(rest)
            if n < 0:
                return None
            data = f.read(n + 2)
            if len(data) != n + 2:
                raise RespError("connection closed inside a bulk string")
            return data[:-2]
        if kind == b"*":
            n = int(rest)
            return None if n < 0 else [cls._read(f) for _ in range(n)]
        raise RespError(f"unexpected reply type {kind!r}")

    def _roundtrip(self, conn, cmds):
        sock, f = conn
        sock.sendall(b"".join(self._encode(c) for c in cmds))
        return [self._read(f) for _ in cmds]

    def pipeline(self, cmds):
        """send every command, then read every reply (error replies come back as RespError objects)"""
        try:
            return self._roundtrip(self._conn(), cmds)
        except (OSError, RespError):
            self.close()  # one reconnect: the store may have dropped an idle connection
            try:
                return self._roundtrip(self._conn(), cmds)
            except OSError as e:
                self.close()
                raise RespError(f"store unreachable at {self.host}:{self.port}: {e}") from e

    def command(self, *args):
        r = self.pipeline([args])[0]
        if isinstance(r, RespError):
            raise r
        return r

    def transaction(self, cmds):
        """MULTI, the commands, EXEC in one write: -> the replies of EXEC; raises if the transaction was not run"""
        replies = self.pipeline([("MULTI",)] + list(cmds) + [("EXEC",)])
        for r in replies[:-1]:
            if isinstance(r, RespError):
                raise r
        res = replies[-1]
        if res is None or isinstance(res, RespError):
            raise RespError("Unable to complete redis transaction." if res is None else str(res))
        for r in res:
            if isinstance(r, RespError):
                raise r
        return res


class RedisTaskStore:
    """`LocalTaskStore`'s contract with the state in a Redis-protocol store instead of this process."""

    def __init__(self, host: str = "127.0.0.1", port: int = 6379, publisher=None, client: RespClient | None = None, **kw):
        self.client = client if client is not None else RespClient(host, port, **kw)
        self.publisher = publisher
        self._paths: dict[str, str] = {}
        self.upserts = 0
        self.gets = 0
        self.round_trips = 0

    # ---- the two HTTP functions, as (status_code, text)
    def upsert(self, body_text: str | None):
        if not body_text or not body_text.strip():
            self.upserts += 1
            return 400, ""
        try:
            task = json.loads(body_text)
            if isinstance(task, list):
                task = task[0]
        except (ValueError, IndexError):
            self.upserts += 1
            return 500, ""
        code, stored = self.upsert_task(task)
        return code, (json.dumps(stored) if code == 200 else "")

    def upsert_task(self, task):
        self.upserts += 1
        prepared = prepare_upsert(task, self._paths)
        if prepared is None:
            return 500, None
        task, task_body, now = prepared
        path, status, tid = task["EndpointPath"], task["BackendStatus"], task["TaskId"]
        publish = bool(task.get("PublishToGrid"))
        cmds = [("SET", tid, json.dumps(task)), ("ZADD", f"{path}_{status}", now, tid)]
        if status == "running":
            cmds.append(("ZREM", f"{path}_created", tid))
        elif status in ("completed", "failed"):
            cmds.append(("ZREM", f"{path}_running", tid))
        if publish and task_body:
            cmds.append(("SET", tid + ORIG_SUFFIX, task_body if isinstance(task_body, str) else json.dumps(task_body)))
        try:
            self.client.transaction(cmds)
            self.round_trips += 1
            if publish and not task_body:  # a later publish of the same task: the original body is in the store (:175)
                raw = self.client.command("GET", tid + ORIG_SUFFIX)
                self.round_trips += 1
                task_body = raw.decode() if raw is not None else None
        except RespError:
            return 500, None  # upstream: any failure inside the try answers 500 (:204-210)
        if publish and self.publisher is not None:
            ok = False
            try:
                ok = bool(self.publisher(dict(task), task_body))
            except Exception:  # noqa: BLE001 - a failed publish moves the task to failed, as upstream (:180-199)
                ok = False
            if not ok:
                failed = dict(task)
                failed["Status"] = "Failed - unable to send to backend service."
                failed["BackendStatus"] = "failed"
                try:
                    self.client.transaction([("SET", tid, json.dumps(failed)), ("ZADD", f"{path}_failed", now, tid),
                                             ("ZREM", f"{path}_created", tid)])
                    self.round_trips += 1
                except RespError:
                    return 500, None
        return 200, task

    def get(self, task_id: str):
        self.gets += 1
        try:
            raw = self.client.command("GET", task_id)
            self.round_trips += 1
        except RespError:
            return 500, ""  # CacheConnectorGet.cs:69-73
        return (200, raw.decode()) if raw is not None else (204, "")

    def get_task(self, task_id: str):
        code, text = self.get(task_id)
        return json.loads(text) if code == 200 else None

    # ---- helpers shared with LocalTaskStore
    def create_task(self, endpoint: str, body: str = "", publish: bool = False) -> dict:
        code, task = self.upsert_task({"TaskId": "", "Status": "created", "BackendStatus": "created",
                                       "Endpoint": endpoint, "Body": body, "PublishToGrid": publish})
        assert code == 200
        return dict(task)

    def set_members(self, key: str):
        flat = self.client.command("ZRANGE", key, 0, -1, "WITHSCORES") or []
        return {flat[i].decode(): int(float(flat[i + 1])) for i in range(0, len(flat), 2)}

    def queue_length(self, key: str) -> int:
        """what Libraries/QueueLogger.cs:44 reads for the autoscaler"""
        return int(self.client.command("ZCARD", key))
