"""Object collectives over an optional process group (reference pg_wrapper.py:15-89).

`pg=None` with torch.distributed initialised means the world group; `pg=None` without it means a single
process, and every collective degenerates to a local assignment.  Only metadata travels through these
(replicated-write assignment, manifest gather, IPC handles); tensor payload never does.
"""

import os
import pickle
from typing import Any, Dict, List, Optional

import numpy as np
import torch
import torch.distributed as dist


def _resolve_group(pg: Optional[Any]) -> Optional[Any]:
    if pg is not None:
        return pg
    if dist.is_available() and dist.is_initialized():
        return dist.group.WORLD
    return None


_SLOT_BYTES = 4096  # per rank; 8 bytes of it carry the pickle's length


def _slots_enabled() -> bool:
    return os.environ.get("TSB200_SLOT_GATHER", "1").lower() not in ("0", "false")


class _SlotGather:
    """Buffers of one process group for the fixed-slot all-gather: a send slot and `world` receive slots, on the
    device for NCCL (with pinned host mirrors) and on the host for gloo."""

    def __init__(self, pg: Any) -> None:
        self.pg = pg  # (kept: the state is looked up by id(pg))
        self.world = dist.get_world_size(group=pg)
        self.nccl = dist.get_backend(group=pg) == "nccl"
        pin = self.nccl and torch.cuda.is_available()
        self.host_send = torch.zeros(_SLOT_BYTES, dtype=torch.uint8, pin_memory=pin)
        self.host_recv = torch.zeros(self.world * _SLOT_BYTES, dtype=torch.uint8, pin_memory=pin)
        self.device: Optional[int] = None
        self.dev_send = self.dev_recv = None
        if self.nccl:
            self.device = torch.cuda.current_device()
            dev = torch.device("cuda", self.device)
            self.dev_send = torch.zeros(_SLOT_BYTES, dtype=torch.uint8, device=dev)
            self.dev_recv = torch.zeros(self.world * _SLOT_BYTES, dtype=torch.uint8, device=dev)
        self.send_np = self.host_send.numpy()
        self.recv_np = self.host_recv.numpy().reshape(self.world, _SLOT_BYTES)


_slot_states: Dict[int, _SlotGather] = {}


def _slot_gather(pg: Any, obj: Any) -> Optional[List[Any]]:
    """Every rank's `obj`, or None when some rank's pickle is longer than a slot (then every rank gets None)."""
    state = _slot_states.get(id(pg))
    if state is None or state.pg is not pg or (state.nccl and state.device != torch.cuda.current_device()):
        state = _slot_states[id(pg)] = _SlotGather(pg)
    data = pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)
    n = len(data)
    state.send_np[:8] = np.frombuffer(np.int64(n).tobytes(), dtype=np.uint8)
    if n <= _SLOT_BYTES - 8:
        state.send_np[8:8 + n] = np.frombuffer(data, dtype=np.uint8)
    if state.nccl:
        state.dev_send.copy_(state.host_send, non_blocking=True)
        dist.all_gather_into_tensor(state.dev_recv, state.dev_send, group=pg)
        state.host_recv.copy_(state.dev_recv, non_blocking=True)
        torch.cuda.current_stream(state.device).synchronize()
    else:
        dist.all_gather(list(state.host_recv.view(state.world, _SLOT_BYTES).unbind(0)), state.host_send, group=pg)
    lengths = state.recv_np[:, :8].copy().view(np.int64).reshape(-1)
    if (lengths > _SLOT_BYTES - 8).any() or (lengths < 0).any():
        return None
    return [pickle.loads(state.recv_np[r, 8:8 + int(lengths[r])].tobytes()) for r in range(state.world)]


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

    def all_gather_object(self, obj_list: List[Any], obj: Any, large: bool = False) -> None:
        """obj_list[r] <- rank r's `obj`.  Small objects (the usual case here: paths, digests, throughputs, IPC
        handles) travel in ONE collective over fixed slots (`_SlotGather`); when any rank's pickle does not fit,
        every rank sees that in the gathered lengths and all of them repeat the exchange with
        dist.all_gather_object (a size exchange, a padded gather and the copies around them).  `large` says the
        caller expects big objects (a manifest): the slots are not tried.  It must be the same on every rank."""
        if self.pg is None:
            obj_list[0] = obj
            return
        if not large and _slots_enabled():
            gathered = _slot_gather(self.pg, obj)
            if gathered is not None:
                obj_list[:] = gathered
                return
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
