"""A minimal flask_restful-compatible `Api` / `Resource` (see compat/__init__.py)."""
from __future__ import annotations

from werkzeug.exceptions import MethodNotAllowed


class Resource:
    methods = ["GET", "POST", "PUT", "DELETE", "PATCH"]


class Api:
    def __init__(self, app=None):
        self.app = app

    def add_resource(self, resource_cls, *urls, resource_class_kwargs=None, resource_class_args=(), endpoint=None, **_):
        kwargs = resource_class_kwargs or {}
        endpoint = endpoint or resource_cls.__name__.lower()
        methods = [m for m in Resource.methods if hasattr(resource_cls, m.lower())]

        def view(**values):
            try:
                from flask import request
            except ImportError:
                from .miniflask import request
            inst = resource_cls(*resource_class_args, **kwargs)
            fn = getattr(inst, request.method.lower(), None)
            if fn is None:
                raise MethodNotAllowed()
            return fn(**values)

        view.__name__ = endpoint
        for url in urls:
            self.app.add_url_rule(url, endpoint=endpoint, view_func=view, methods=methods)
