"""A small in-process Redis-protocol (RESP2) server for the task-store tests: strings, sorted sets and MULTI / EXEC,
the commands the platform's cache-connector functions use (ProcessManager/CacheManager/CacheConnectorUpsert.cs:123-163:
StringSet, SortedSetAdd, SortedSetRemove in one transaction; CacheConnectorGet.cs:56 StringGet; Libraries/QueueLogger.cs:44
SortedSetLength).  Test infrastructure: no Redis server binary exists in the image, and the client under test
(`task_management/redis_task_store.py`) must be exercised over a real socket."""
import socketserver
import threading


class _State:
    def __init__(self):
        self.lock = threading.Lock()
        self.strings = {}
        self.zsets = {}
        self.commands = []   # every command received, upper-cased name first (tests look at the wire traffic)
        self.fail_exec = 0   # answer nil to this many EXECs (an aborted transaction)


def _bulk(b):
    return b"$-1\r\n" if b is None else b"$%d\r\n%s\r\n" % (len(b), b)


def _num(x):
    s = repr(float(x))
    return (s[:-2] if s.endswith(".0") else s).encode()


def _apply(st, cmd):
    """one command against the data (caller holds the lock) -> encoded reply"""
    name, a = cmd[0].upper(), cmd[1:]
    if name == b"PING":
        return b"+PONG\r\n"
    if name == b"SET" and len(a) == 2:
        if a[0] in st.zsets:
            return b"-WRONGTYPE Operation against a key holding the wrong kind of value\r\n"
        st.strings[a[0]] = a[1]
        return b"+OK\r\n"
    if name == b"GET" and len(a) == 1:
        if a[0] in st.zsets:
            return b"-WRONGTYPE Operation against a key holding the wrong kind of value\r\n"
        return _bulk(st.strings.get(a[0]))
    if name == b"DEL":
        n = 0
        for k in a:
            n += (st.strings.pop(k, None) is not None) + (st.zsets.pop(k, None) is not None)
        return b":%d\r\n" % n
    if name == b"ZADD" and len(a) >= 3 and len(a) % 2 == 1:
        if a[0] in st.strings:
            return b"-WRONGTYPE Operation against a key holding the wrong kind of value\r\n"
        z, added = st.zsets.setdefault(a[0], {}), 0
        try:
            pairs = [(float(a[i]), a[i + 1]) for i in range(1, len(a), 2)]
        except ValueError:
            return b"-ERR value is not a valid float\r\n"
        for score, member in pairs:
            added += member not in z
            z[member] = score
        return b":%d\r\n" % added
    if name == b"ZREM" and len(a) >= 2:
        z, n = st.zsets.get(a[0], {}), 0
        for m in a[1:]:
            n += z.pop(m, None) is not None
        if not z:
            st.zsets.pop(a[0], None)
        return b":%d\r\n" % n
    if name == b"ZCARD" and len(a) == 1:
        return b":%d\r\n" % len(st.zsets.get(a[0], {}))
    if name == b"ZSCORE" and len(a) == 2:
        s = st.zsets.get(a[0], {}).get(a[1])
        return _bulk(None if s is None else _num(s))
    if name == b"ZRANGE" and len(a) >= 3:
        items = sorted(st.zsets.get(a[0], {}).items(), key=lambda kv: (kv[1], kv[0]))
        lo, hi = int(a[1]), int(a[2])
        n = len(items)
        lo = max(n + lo, 0) if lo < 0 else lo
        hi = n + hi if hi < 0 else min(hi, n - 1)
        sel = items[lo:hi + 1] if lo <= hi else []
        withscores = len(a) > 3 and a[3].upper() == b"WITHSCORES"
        out = [b"*%d\r\n" % (len(sel) * (2 if withscores else 1))]
        for m, s in sel:
            out.append(_bulk(m))
            if withscores:
                out.append(_bulk(_num(s)))
        return b"".join(out)
    if name == b"DBSIZE":
        return b":%d\r\n" % (len(st.strings) + len(st.zsets))
    if name == b"FLUSHALL":
        st.strings.clear()
        st.zsets.clear()
        return b"+OK\r\n"
    if name in (b"SELECT", b"AUTH"):
        return b"+OK\r\n"
    return b"-ERR unknown command or wrong number of arguments for '%s'\r\n" % cmd[0]


class _Handler(socketserver.StreamRequestHandler):
    def _read_command(self):
        line = self.rfile.readline()
        if not line:
            return None
        if not line.startswith(b"*"):
            return line.split()  # inline command
        n = int(line[1:])
        args = []
        for _ in range(n):
            hdr = self.rfile.readline()
            size = int(hdr[1:])
            args.append(self.rfile.read(size + 2)[:-2])
        return args

    def handle(self):
        st = self.server.state
        queued = None
        while True:
            try:
                cmd = self._read_command()
            except (ValueError, OSError):
                return
            if not cmd:
                return
            name = cmd[0].upper()
            with st.lock:
                st.commands.append([name] + cmd[1:])
            if name == b"MULTI":
                queued = []
                reply = b"+OK\r\n"
            elif name == b"DISCARD":
                queued = None
                reply = b"+OK\r\n"
            elif name == b"EXEC":
                if queued is None:
                    reply = b"-ERR EXEC without MULTI\r\n"
                else:
                    with st.lock:  # all or nothing, nobody in between
                        if st.fail_exec > 0:
                            st.fail_exec -= 1
                            reply = b"*-1\r\n"
                        else:
                            parts = [_apply(st, c) for c in queued]
                            reply = b"*%d\r\n" % len(parts) + b"".join(parts)
                    queued = None
            elif queued is not None:
                queued.append(cmd)
                reply = b"+QUEUED\r\n"
            else:
                with st.lock:
                    reply = _apply(st, cmd)
            try:
                self.wfile.write(reply)
                self.wfile.flush()
            except OSError:
                return


class FakeRedis:
    """`with FakeRedis() as r:` -> r.host, r.port, r.state"""

    def __init__(self, host="127.0.0.1", port=0):
        socketserver.ThreadingTCPServer.allow_reuse_address = True
        self._server = socketserver.ThreadingTCPServer((host, port), _Handler)
        self._server.daemon_threads = True
        self.state = self._server.state = _State()
        self.host, self.port = self._server.server_address[:2]
        self._thread = threading.Thread(target=self._server.serve_forever, daemon=True, name="fake-redis")
        self._thread.start()

    def close(self):
        self._server.shutdown()
        self._thread.join(timeout=5)
        self._server.server_close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
