"""`TaskManager`: the facade that `APIService` and user task functions call.

Mirrors the public methods of the reference class
(/root/reference/APIs/1.0/base-py/task_management/api_task.py:8-38): AddTask, CompleteTask,
FailTask, UpdateTaskStatus, AddPipelineTask, GetTaskStatus, each answered by the
distributed task client.  Tasks are created by the platform; a container only adopts the task
whose id arrives in the `taskId` request header (reference :12-20).
"""
import logging

from task_management.distributed_api_task import DistributedApiTaskManager

try:
    from flask_restful import Resource
except ImportError:  # the build image has no flask_restful
    from ai4e_b200.compat.minirestful import Resource

_log = logging.getLogger("ai4e.task_manager")


def _header_task_id(request):
    """the platform-assigned task id carried by the request, or None."""
    headers = getattr(request, "headers", None) if request else None
    if headers is None or "taskId" not in headers:
        return None
    return headers.get("taskId")


class TaskManager:
    def __init__(self, store=None):
        self.distributed_api_task = DistributedApiTaskManager(store)

    def AddTask(self, request):
        task_id = _header_task_id(request)
        backend = self.distributed_api_task
        return backend.AddTask() if task_id is None else backend.GetTaskStatus(task_id)

    def GetTaskStatus(self, taskId):
        return self.distributed_api_task.GetTaskStatus(taskId)

    def UpdateTaskStatus(self, taskId, status):
        _log.debug("task %s -> running: %s", taskId, status)
        return self.distributed_api_task.UpdateTaskStatus(taskId, status)

    def CompleteTask(self, taskId, status):
        _log.debug("task %s -> completed: %s", taskId, status)
        return self.distributed_api_task.CompleteTask(taskId, status)

    def FailTask(self, taskId, status):
        _log.debug("task %s -> failed: %s", taskId, status)
        self.distributed_api_task.FailTask(taskId, status)  # like upstream, returns nothing

    def AddPipelineTask(self, taskId, organization_moniker, version, api_name, body):
        _log.debug("task %s -> pipeline stage %s/%s/%s", taskId, version, organization_moniker, api_name)
        return self.distributed_api_task.AddPipelineTask(taskId, organization_moniker, version, api_name, body)


class Task(Resource):
    """REST resource over the task client (the reference ships one here too, :40-46)."""

    def __init__(self, **kwargs):
        self.distributed_api_task = kwargs["distributed_api_task"]

    def get(self, id):
        return self.distributed_api_task.GetTaskStatus(id)
