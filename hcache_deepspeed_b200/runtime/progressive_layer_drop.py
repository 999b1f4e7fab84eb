"""Progressive Layer Drop: keep probability theta(t) = (1 - theta) * exp(-gamma * t) + theta, handed to the model's
forward as ``progressive_layer_drop=True, pld_theta=...``. Reference: ``deepspeed/runtime/progressive_layer_drop.py``."""
import math


class ProgressiveLayerDrop:

    def __init__(self, theta: float = 0.5, gamma: float = 0.001):
        self.theta, self.gamma = theta, gamma
        self.current_theta = 1.0

    def get_state(self):
        return {"progressive_layer_drop": True, "pld_theta": self.get_theta()}

    def get_theta(self) -> float:
        return self.current_theta

    def update_state(self, global_step: int):
        self.current_theta = (1.0 - self.theta) * math.exp(-self.gamma * global_step) + self.theta
