"""Object collectives over an optional process group (reference pg_wrapper.py:15-89).

`pg=None` with torch.distributed initialised means the world group; `pg=None` without it means a single
process, and every collective degenerates to a local assignment.  Only metadata travels through these
(replicated-write assignment, manifest gather, IPC handles); tensor payload never does.
"""

from typing import Any, List, Optional

import torch
import torch.distributed as dist


def _resolve_group(pg: Optional[Any]) -> Optional[Any]:
    if pg is not None:
        return pg
    if dist.is_available() and dist.is_initialized():
        return dist.group.WORLD
    return None


class PGWrapper:
    def __init__(self, pg: Optional[Any] = None) -> None:
        self.pg: Optional[Any] = _resolve_group(pg)

    # -- topology --------------------------------------------------------------------------------
    def get_rank(self) -> int:
        return dist.get_rank(group=self.pg) if self.pg is not None else 0

    def get_world_size(self) -> int:
        return dist.get_world_size(group=self.pg) if self.pg is not None else 1

    def _is_nccl(self) -> bool:
        return self.pg is not None and dist.get_backend(group=self.pg) == "nccl"

    def _global_rank(self, group_rank: int) -> int:
        return dist.get_global_rank(self.pg, group_rank)

    # -- collectives -----------------------------------------------------------------------------
    def barrier(self) -> None:
        if self.pg is None:
            return
        if self._is_nccl():  # name the device: NCCL otherwise guesses it from the rank
            dist.barrier(group=self.pg, device_ids=[torch.cuda.current_device()])
        else:
            dist.barrier(group=self.pg)

    def all_gather_object(self, obj_list: List[Any], obj: Any) -> None:
        if self.pg is None:
            obj_list[0] = obj
        else:
            dist.all_gather_object(obj_list, obj, group=self.pg)

    def broadcast_object_list(self, obj_list: List[Any], src: int = 0) -> None:
        if self.pg is not None:
            dist.broadcast_object_list(obj_list, src=self._global_rank(src), group=self.pg)

    def scatter_object_list(self, output_list: List[Any], input_list: Optional[List[Any]], src: int = 0) -> None:
        """output_list[0] <- element `rank` of the list rank `src` supplies."""
        me, world = self.get_rank(), self.get_world_size()
        if me == src:
            if input_list is None:
                raise RuntimeError("scatter_object_list: the source rank has to supply input_list.")
            if len(input_list) != world:
                raise RuntimeError(
                    f"scatter_object_list: input_list has {len(input_list)} elements for a group of {world} ranks."
                )
            outgoing = input_list
        else:
            outgoing = [None] * world
        if self.pg is None:
            output_list[0] = outgoing[0]
        elif self._is_nccl():
            # NCCL has no object scatter: everybody receives the list and keeps its own element
            self.broadcast_object_list(outgoing, src=src)
            output_list[0] = outgoing[me]
        else:
            dist.scatter_object_list(output_list, outgoing, src=self._global_rank(src), group=self.pg)
