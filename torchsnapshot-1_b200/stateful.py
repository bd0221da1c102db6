"""What `Snapshot` accepts as application state: a dict of objects that can export and re-import their state.
Mirrors the reference's protocol (torchsnapshot/stateful.py:14-25) so `isinstance(x, Stateful)` means the same.
"""

from typing import Any, Dict, Protocol, TypeVar, runtime_checkable

_State = Dict[str, Any]


@runtime_checkable
class Stateful(Protocol):
    """Modules, optimizers, LR schedulers, `StateDict`, `RNGState`, ... -- anything with this pair."""

    def state_dict(self) -> _State:
        """The state as a (nested) container of tensors, primitives and picklable objects."""

    def load_state_dict(self, state_dict: _State) -> None:
        """Take back what `state_dict()` produced."""


T = TypeVar("T", bound=Stateful)
AppState = Dict[str, T]  # key = the stateful's name: first component of every logical path in the manifest
