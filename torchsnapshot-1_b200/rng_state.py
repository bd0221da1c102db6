"""RNGState: captures torch's global RNG state as a Stateful (reference rng_state.py:13-38)."""

from typing import Dict

import torch


class RNGState:
    def state_dict(self) -> Dict[str, torch.Tensor]:
        return {"rng_state": torch.get_rng_state()}

    def load_state_dict(self, state_dict: Dict[str, torch.Tensor]) -> None:
        torch.set_rng_state(state_dict["rng_state"])
