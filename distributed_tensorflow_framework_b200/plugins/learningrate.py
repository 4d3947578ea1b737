"""Learning-rate persistence. Parity: /root/reference/distribute_learningrate.py:12-50.

JSON file `{"learning_rate": x}`; loaded at construction if the file exists. Fixes: `load()` no longer touches an
unset attribute when the file vanished; `decay_factor`, stored but never used by the reference
(:19), drives the optional `decay()` helper.
"""
import json
import os

from ..api import log as logger


class LearningRate(object):
    def __init__(self, initial_learning_rate, save_path, decay_factor=None):
        self.save_path = save_path
        self.learning_rate = initial_learning_rate
        self.decay_factor = decay_factor
        if save_path and os.path.exists(save_path):
            self.learning_rate = self.load()

    def save(self, current_learning_rate=None):
        if current_learning_rate is not None:
            self.learning_rate = current_learning_rate
        if not os.path.exists(self.save_path):
            logger.info("Create Json file for learning rate.")
        tmp = self.save_path + ".tmp"
        with open(tmp, "w") as f:
            json.dump({"learning_rate": self.learning_rate}, f)
        os.replace(tmp, self.save_path)

    def load(self):
        if not os.path.exists(self.save_path):
            return self.learning_rate
        with open(self.save_path, "r") as f:
            data = json.load(f)
        self.learning_rate = data["learning_rate"]
        return self.learning_rate

    def update(self, current_learning_rate):
        self.learning_rate = current_learning_rate

    def decay(self):
        if self.decay_factor is not None:
            self.learning_rate = self.learning_rate * self.decay_factor
        return self.learning_rate
