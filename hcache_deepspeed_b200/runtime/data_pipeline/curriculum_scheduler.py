"""Curriculum-learning difficulty schedules: fixed_discrete, fixed_linear, fixed_root, custom.
Reference: ``deepspeed/runtime/data_pipeline/curriculum_scheduler.py:11-158``."""
from __future__ import annotations

import math

CURRICULUM_LEARNING_MIN_DIFFICULTY, CURRICULUM_LEARNING_MAX_DIFFICULTY = "min_difficulty", "max_difficulty"
CURRICULUM_LEARNING_SCHEDULE_TYPE, CURRICULUM_LEARNING_SCHEDULE_CONFIG = "schedule_type", "schedule_config"
FIXED_DISCRETE, FIXED_LINEAR, FIXED_ROOT, CUSTOM = "fixed_discrete", "fixed_linear", "fixed_root", "custom"


class CurriculumScheduler:

    def __init__(self, config: dict):
        for k in (CURRICULUM_LEARNING_MIN_DIFFICULTY, CURRICULUM_LEARNING_MAX_DIFFICULTY, CURRICULUM_LEARNING_SCHEDULE_TYPE):
            assert k in config, f"Curriculum learning requires the config '{k}'"
        self.state = {"min_difficulty": config["min_difficulty"], "max_difficulty": config["max_difficulty"],
                      "current_difficulty": config["min_difficulty"], "schedule_type": config["schedule_type"],
                      "schedule": {}}
        self.first_step = True
        self.custom_get_difficulty = None
        st, sc = config["schedule_type"], config.get("schedule_config", {})
        if st == FIXED_DISCRETE:
            assert "difficulty" in sc and "max_step" in sc and len(sc["max_step"]) > 0
            assert len(sc["difficulty"]) == len(sc["max_step"]) + 1
            self.state["schedule"] = sc
        elif st in (FIXED_ROOT, FIXED_LINEAR):
            assert "total_curriculum_step" in sc and "difficulty_step" in sc
            if st == FIXED_ROOT:
                assert "root_degree" in sc
            self.state["schedule"] = sc
        elif st == CUSTOM:
            pass
        else:
            raise RuntimeError("Unsupported curriculum schedule type")

    def get_current_difficulty(self):
        return self.state["current_difficulty"]

    def set_current_difficulty(self, d):
        self.state["current_difficulty"] = d

    def set_custom_get_difficulty(self, fn):
        self.custom_get_difficulty = fn

    def get_state(self):
        return self.state

    def set_state(self, state):
        self.state = state

    def _fixed_discrete(self, step):
        s = self.state["schedule"]
        if step > s["max_step"][-1]:
            return s["difficulty"][-1]
        for i, ms in enumerate(s["max_step"]):
            if step <= ms:
                return s["difficulty"][i]

    def _fixed_root(self, step, root_degree=None):
        s = self.state["schedule"]
        root_degree = s["root_degree"] if root_degree is None else root_degree
        frac = (float(step) / s["total_curriculum_step"]) ** (1.0 / root_degree)
        d = math.floor(frac * (self.state["max_difficulty"] - self.state["min_difficulty"]) + self.state["min_difficulty"])
        d -= d % s["difficulty_step"]
        return min(max(d, self.state["min_difficulty"]) if d > 0 else self.state["min_difficulty"], self.state["max_difficulty"])

    def get_difficulty(self, global_steps):
        st = self.state["schedule_type"]
        if st == FIXED_DISCRETE:
            return self._fixed_discrete(global_steps)
        if st == FIXED_LINEAR:
            return self._fixed_root(global_steps, 1)
        if st == FIXED_ROOT:
            return self._fixed_root(global_steps)
        if st == CUSTOM:
            return self.custom_get_difficulty(global_steps)
        raise RuntimeError("Unsupported curriculum schedule type")

    def update_difficulty(self, global_steps):
        if self.state["current_difficulty"] < self.state["max_difficulty"]:
            self.state["current_difficulty"] = self.get_difficulty(global_steps)
        return self.state["current_difficulty"]
