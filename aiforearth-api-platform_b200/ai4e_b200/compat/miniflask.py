"""A minimal Flask-compatible application object on top of Werkzeug: exactly the surface
`ai4e_service.APIService` and typical AI4E API projects touch (`Flask`, `request`, `abort`,
`current_app`, `add_url_rule`, `route`, `before_request`, `test_client`, `run`)."""
from __future__ import annotations

import contextvars
import json

from werkzeug.exceptions import HTTPException, InternalServerError, MethodNotAllowed, NotFound
from werkzeug.exceptions import abort as _wz_abort
from werkzeug.local import LocalProxy
from werkzeug.routing import Map, Rule
from werkzeug.test import Client
from werkzeug.wrappers import Request, Response

_request_var: contextvars.ContextVar = contextvars.ContextVar("miniflask_request")
_app_var: contextvars.ContextVar = contextvars.ContextVar("miniflask_app")

request = LocalProxy(_request_var.get)
current_app = LocalProxy(_app_var.get)


def abort(code, *args, **kwargs):
    """flask.abort: raises the HTTPException for `code`; a second argument is the description."""
    _wz_abort(code, *args, **kwargs)


def jsonify(obj):
    return Response(json.dumps(obj), mimetype="application/json")


class _Views:
    pass


views = _Views()


class Flask:
    def __init__(self, import_name="app"):
        self.import_name = import_name
        self.url_map = Map()
        self.view_functions = {}
        self.before_request_funcs = []
        self.config = {}
        self._static_rules = {}  # path without converters -> (endpoint, methods): matched by one dict look-up

    # ---- registration
    def add_url_rule(self, rule, endpoint=None, view_func=None, methods=None, provide_automatic_options=None, **_):
        endpoint = endpoint or (view_func.__name__ if view_func else rule)
        methods = [m.upper() for m in (methods or ["GET"])]
        if provide_automatic_options or provide_automatic_options is None:
            if "OPTIONS" not in methods:
                methods = methods + ["OPTIONS"]
        if endpoint in self.view_functions and self.view_functions[endpoint] is not view_func:
            raise AssertionError(f"View function mapping is overwriting an existing endpoint function: {endpoint}")
        self.url_map.add(Rule(rule, endpoint=endpoint, methods=methods))
        self.view_functions[endpoint] = view_func
        if "<" not in rule and rule not in self._static_rules:
            self._static_rules[rule] = (endpoint, frozenset(methods))

    def route(self, rule, **options):
        def deco(f):
            self.add_url_rule(rule, options.pop("endpoint", None), f, **options)
            return f
        return deco

    def before_request(self, f):
        self.before_request_funcs.append(f)
        return f

    # ---- dispatch
    def make_response(self, rv):
        if isinstance(rv, Response):
            return rv
        if isinstance(rv, HTTPException):
            return rv.get_response()
        status = None
        if isinstance(rv, tuple):
            rv, status = rv[0], rv[1]
        if isinstance(rv, (dict, list)):
            resp = jsonify(rv)
        elif isinstance(rv, bytes):
            resp = Response(rv)
        elif rv is None:
            raise TypeError("The view function did not return a valid response")
        else:
            resp = Response(str(rv), mimetype="text/html")
        if status is not None:
            resp.status_code = int(status)
        return resp

    def full_dispatch(self, req: Request) -> Response:
        try:
            for f in self.before_request_funcs:
                rv = f()
                if rv is not None:
                    return self.make_response(rv)
            hit = self._static_rules.get(req.path)
            if hit is not None and req.method in hit[1] and req.method != "OPTIONS":
                return self.make_response(self.view_functions[hit[0]]())  # a fixed path: no need for the matcher
            adapter = self.url_map.bind_to_environ(req.environ)
            endpoint, values = adapter.match()
            if req.method == "OPTIONS":
                resp = Response()
                resp.headers["Allow"] = ", ".join(sorted(adapter.allowed_methods()))
                return resp
            return self.make_response(self.view_functions[endpoint](**values))
        except HTTPException as e:
            return e.get_response(req.environ)
        except Exception:  # noqa: BLE001 - a web server answers 500
            if self.config.get("PROPAGATE_EXCEPTIONS") or self.config.get("TESTING"):
                raise
            return InternalServerError().get_response(req.environ)

    def wsgi_app(self, environ, start_response):
        req = Request(environ)
        t1, t2 = _request_var.set(req), _app_var.set(self)
        try:
            return self.full_dispatch(req)(environ, start_response)
        finally:
            _request_var.reset(t1)
            _app_var.reset(t2)

    def __call__(self, environ, start_response):
        return self.wsgi_app(environ, start_response)

    def test_client(self):
        return Client(self, Response)

    def run(self, host="127.0.0.1", port=5000, threaded=True, **_):  # pragma: no cover
        from werkzeug.serving import run_simple
        run_simple(host, port, self, threaded=threaded)


__all__ = ["Flask", "request", "abort", "current_app", "views", "jsonify", "Response", "NotFound", "MethodNotAllowed"]
