"""Uniform object collectives over an optional process group (reference pg_wrapper.py:15-89).
Metadata-sized messages only: replicated-write assignment and manifest gather; tensor payload
never travels through these."""

from typing import Any, List, Optional

import torch
import torch.distributed as dist


class PGWrapper:
    def __init__(self, pg: Optional[Any] = None) -> None:
        if pg is None and dist.is_available() and dist.is_initialized():
            self.pg: Optional[Any] = dist.group.WORLD
        else:
            self.pg = pg

    def get_rank(self) -> int:
        return 0 if self.pg is None else dist.get_rank(group=self.pg)

    def get_world_size(self) -> int:
        return 1 if self.pg is None else dist.get_world_size(group=self.pg)

    def barrier(self) -> None:
        if self.pg is None:
            return
        if dist.get_backend(group=self.pg) == "nccl":
            dist.barrier(group=self.pg, device_ids=[torch.cuda.current_device()])
        else:
            dist.barrier(group=self.pg)

    def all_gather_object(self, obj_list: List[Any], obj: Any) -> None:
        if self.pg is None:
            obj_list[0] = obj
            return
        dist.all_gather_object(obj_list, obj, group=self.pg)

    def broadcast_object_list(self, obj_list: List[Any], src: int = 0) -> None:
        if self.pg is None:
            return
        dist.broadcast_object_list(obj_list, src=dist.get_global_rank(self.pg, src), group=self.pg)
