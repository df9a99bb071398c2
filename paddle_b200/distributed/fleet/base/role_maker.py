"""Role makers, collective subset (reference: python/paddle/distributed/fleet/base/role_maker.py)."""
from __future__ import annotations

import os


class Role:
    WORKER = 1
    SERVER = 2
    HETER_WORKER = 3
    ALL = 4
    COORDINATOR = 5


class RoleMakerBase:
    def __init__(self):
        self._role = Role.WORKER
        self._current_id = 0
        self._worker_endpoints = []

    def _is_worker(self):
        return True

    def _is_server(self):
        return False

    def _is_first_worker(self):
        return self._worker_index() == 0

    def _worker_num(self):
        return max(len(self._worker_endpoints), 1)

    def _worker_index(self):
        return self._current_id

    def _get_trainer_endpoints(self):
        return self._worker_endpoints

    def _is_collective(self):
        return True


class PaddleCloudRoleMaker(RoleMakerBase):
    def __init__(self, is_collective=False, **kwargs):
        super().__init__()
        self._is_collective_flag = is_collective
        self._kwargs = kwargs
        self._generated = False

    def _generate_role(self):
        if self._generated:
            return
        e = os.environ
        self._current_id = int(e.get("PADDLE_TRAINER_ID", e.get("RANK", "0")))
        eps = e.get("PADDLE_TRAINER_ENDPOINTS", "")
        self._worker_endpoints = eps.split(",") if eps else ["127.0.0.1:6170"] * int(e.get("PADDLE_TRAINERS_NUM", e.get("WORLD_SIZE", "1")))
        self._cur_endpoint = e.get("PADDLE_CURRENT_ENDPOINT", "")
        self._trainers_num = int(e.get("PADDLE_TRAINERS_NUM", e.get("WORLD_SIZE", "1")))
        self._generated = True

    def _worker_num(self):
        self._generate_role()
        return self._trainers_num

    def _worker_index(self):
        self._generate_role()
        return self._current_id

    def _is_collective(self):
        return self._is_collective_flag


class UserDefinedRoleMaker(PaddleCloudRoleMaker):
    def __init__(self, is_collective=False, init_gloo=False, **kwargs):
        super().__init__(is_collective, **kwargs)
        self._current_id = kwargs.get("current_id", 0)
        self._worker_endpoints = kwargs.get("worker_endpoints", [])
        self._trainers_num = kwargs.get("worker_num", max(len(self._worker_endpoints), 1))
        self._generated = True
