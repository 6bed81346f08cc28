"""`TaskManager` - the facade user code and `APIService` call; same methods and delegation as
the reference (/root/reference/APIs/1.0/base-py/task_management/api_task.py:8-38)."""
from task_management.distributed_api_task import DistributedApiTaskManager

try:
    from flask_restful import Resource
except ImportError:  # build image has no flask_restful
    from ai4e_b200.compat.minirestful import Resource

print("Creating API task manager.")


class TaskManager:
    def __init__(self, store=None):
        self.distributed_api_task = DistributedApiTaskManager(store)

    def AddTask(self, request):
        # the platform creates tasks; a request that carries a `taskId` header reuses that task
        if not request or 'taskId' not in request.headers:
            return self.distributed_api_task.AddTask()
        return self.distributed_api_task.GetTaskStatus(request.headers.get('taskId'))

    def CompleteTask(self, taskId, status):
        print("Completing task: " + str(taskId) + " to:" + str(status))
        return self.distributed_api_task.CompleteTask(taskId, status)

    def FailTask(self, taskId, status):
        self.distributed_api_task.FailTask(taskId, status)

    def UpdateTaskStatus(self, taskId, status):
        print("Updating task: " + str(taskId) + " to:" + str(status))
        return self.distributed_api_task.UpdateTaskStatus(taskId, status)

    def AddPipelineTask(self, taskId, organization_moniker, version, api_name, body):
        print("Adding pipeline task: " + str(taskId))
        return self.distributed_api_task.AddPipelineTask(taskId, organization_moniker, version, api_name, body)

    def GetTaskStatus(self, taskId):
        return self.distributed_api_task.GetTaskStatus(taskId)


class Task(Resource):
    def __init__(self, **kwargs):
        self.distributed_api_task = kwargs['distributed_api_task']

    def get(self, id):
        return self.distributed_api_task.GetTaskStatus(id)
