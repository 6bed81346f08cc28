"""HTTP facade of the platform's two cache-connector functions over a `LocalTaskStore` (state in this process) or a
`RedisTaskStore` (state in a Redis-protocol store, the reference's keys), so that a
container (this repo's client over `requests`, or the reference's own aiohttp client, unmodified)
can be driven against the real wire contract without Azure:

    POST {upsert}            body = APITask JSON (or a one-element array)  -> 200 + stored JSON | 400 | 500
    GET  {get}?taskId=<id>                                                 -> 200 + stored JSON | 204

Contract restated from the reference (citations relative to /root/reference):
  ProcessManager/CacheManager/CacheConnectorUpsert.cs:40-41 (POST, route = function name), :57-74 (empty body
  -> 400), :60-67 (array body -> first element), :76-82 (undeserialisable body -> 500), :212 (200 + the
  serialised task); ProcessManager/CacheManager/CacheConnectorGet.cs:26-27 (GET), :33-37 (query key
  `taskId`), :56-66 (200 + JSON, 204 when unknown).  The state machine itself is `LocalTaskStore` / `RedisTaskStore`.

    store = LocalTaskStore()
    with serve(store) as srv:                       # a thread with a Werkzeug server on 127.0.0.1
        os.environ['CACHE_CONNECTOR_UPSERT_URI'] = srv.upsert_uri
        os.environ['CACHE_CONNECTOR_GET_URI'] = srv.get_uri
"""
from __future__ import annotations

import threading

from werkzeug.wrappers import Request, Response

from task_management.local_task_store import LocalTaskStore

UPSERT_PATH = "/api/CacheConnectorUpsert"
GET_PATH = "/api/CacheConnectorGet"


class CacheConnectorFacade:
    """WSGI application: the two functions, nothing else."""

    def __init__(self, store: LocalTaskStore | None = None):
        self.store = store if store is not None else LocalTaskStore()
        self.requests = {"upsert": 0, "get": 0}
        self._guard = threading.Lock()

    def __call__(self, environ, start_response):
        req = Request(environ)
        path = req.path.rstrip("/")
        if path.endswith(UPSERT_PATH) or path.endswith("/CacheConnectorUpsert"):
            if req.method != "POST":
                return Response(status=404)(environ, start_response)
            with self._guard:
                self.requests["upsert"] += 1
            code, text = self.store.upsert(req.get_data(as_text=True))
            return Response(text, status=code, mimetype="text/plain")(environ, start_response)
        if path.endswith(GET_PATH) or path.endswith("/CacheConnectorGet"):
            if req.method != "GET":
                return Response(status=404)(environ, start_response)
            with self._guard:
                self.requests["get"] += 1
            code, text = self.store.get(req.args.get("taskId", "nil"))
            return Response(text, status=code, mimetype="text/plain")(environ, start_response)
        return Response(status=404)(environ, start_response)


class _Served:
    def __init__(self, facade, server, thread):
        self.facade, self.store = facade, facade.store
        self._server, self._thread = server, thread
        host, port = server.server_address[:2]
        self.base = f"http://{host}:{port}"
        self.upsert_uri = self.base + UPSERT_PATH
        self.get_uri = self.base + GET_PATH

    def close(self):
        self._server.shutdown()
        self._thread.join(timeout=5)
        self._server.server_close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def serve(store: LocalTaskStore | None = None, host: str = "127.0.0.1", port: int = 0) -> _Served:
    """start the facade on a background thread (port 0 = any free port)."""
    from werkzeug.serving import make_server
    facade = CacheConnectorFacade(store)
    server = make_server(host, port, facade, threaded=True)
    thread = threading.Thread(target=server.serve_forever, daemon=True, name="ai4e-cache-connector")
    thread.start()
    return _Served(facade, server, thread)


def main(argv=None):
    """`python -m task_management.cache_connector_http [--port 7071] [--redis host:port]`: the two functions as a
    process of their own, state in this process or in a Redis-protocol store (the reference's keys)."""
    import argparse
    import time
    ap = argparse.ArgumentParser(description=main.__doc__)
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=7071)
    ap.add_argument("--redis", default=None, help="host:port of a Redis-protocol store; default: in-process state")
    a = ap.parse_args(argv)
    store = None
    if a.redis:
        from task_management.redis_task_store import RedisTaskStore
        host, _, port = a.redis.rpartition(":")
        store = RedisTaskStore(host or "127.0.0.1", int(port))
    with serve(store, a.host, a.port) as srv:
        print(f"CACHE_CONNECTOR_UPSERT_URI={srv.upsert_uri}\nCACHE_CONNECTOR_GET_URI={srv.get_uri}", flush=True)
        try:
            while True:
                time.sleep(3600)
        except KeyboardInterrupt:
            pass


if __name__ == "__main__":
    main()
