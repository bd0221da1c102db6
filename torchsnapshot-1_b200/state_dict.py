"""`StateDict`: a dict that is its own state dict, for state that has no `state_dict()` of its own
(reference torchsnapshot/state_dict.py:13-41 -- same semantics: `state_dict()` hands out the live mapping
and `load_state_dict()` updates it, so keys absent from a snapshot keep their current values)."""

from collections import UserDict
from typing import Any, Dict


class StateDict(UserDict):
    def state_dict(self) -> Dict[str, Any]:
        return self.data

    def load_state_dict(self, state_dict: Dict[str, Any]) -> None:
        self.data.update(state_dict)
