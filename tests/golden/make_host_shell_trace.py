"""Run the scripted scenario against the UNMODIFIED reference host shell
(/root/reference/APIs/1.0/base-py/ai4e_service.py) and freeze what it does into
tests/golden/host_shell_trace.json.

    python tests/golden/make_host_shell_trace.py

The reference imports flask, flask_restful, ai4e_app_insights_wrapper and azure_monitor_logger,
none of which exist in this image, so they are provided as `sys.modules` stubs:
flask / flask_restful by this repo's Werkzeug shims, the two Azure modules by empty classes, and
`task_management.api_task` by a placeholder (the scenario swaps in a recording task manager).
The reference file itself is imported from where it lies; nothing is copied.
"""
import importlib.util
import json
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference/APIs/1.0/base-py/ai4e_service.py"


def load_reference_service():
    sys.path.insert(0, os.path.join(ROOT, "aiforearth-api-platform_b200"))
    sys.path.insert(0, ROOT)
    from ai4e_b200.compat import miniflask, minirestful
    flask = types.ModuleType("flask")
    for n in ("Flask", "abort", "request", "current_app", "views"):
        setattr(flask, n, getattr(miniflask, n))
    restful = types.ModuleType("flask_restful")
    restful.Resource, restful.Api = minirestful.Resource, minirestful.Api
    insights = types.ModuleType("ai4e_app_insights_wrapper")
    insights.AI4EAppInsights = type("AI4EAppInsights", (), {})
    monitor = types.ModuleType("azure_monitor_logger")
    monitor.AzureMonitorLogger = type("AzureMonitorLogger", (), {})
    tm_pkg = types.ModuleType("task_management")
    tm_mod = types.ModuleType("task_management.api_task")
    tm_mod.TaskManager = type("TaskManager", (), {})
    tm_pkg.api_task = tm_mod
    sys.modules.update({"flask": flask, "flask_restful": restful, "ai4e_app_insights_wrapper": insights,
                        "azure_monitor_logger": monitor, "task_management": tm_pkg,
                        "task_management.api_task": tm_mod})
    os.environ["API_PREFIX"] = ""
    spec = importlib.util.spec_from_file_location("reference_ai4e_service", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod, miniflask.Flask


def main():
    mod, Flask = load_reference_service()
    from tests.host_shell_scenario import run_scenario
    trace = run_scenario(mod.APIService, Flask)
    trace["_source"] = REF
    if "--print" in sys.argv:
        print(json.dumps(trace))
        return
    path = os.path.join(ROOT, "tests", "golden", "host_shell_trace.json")
    json.dump(trace, open(path, "w"), indent=1, sort_keys=True)
    print("wrote", path)


if __name__ == "__main__":
    main()
