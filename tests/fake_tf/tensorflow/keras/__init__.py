"""Stand-in for the bits of tf.keras that b200dp.hvd_keras touches (see ../__init__.py)."""
import types

import torch

from .. import Variable


class Callback:
    def __init__(self):
        self.model, self.params = None, {}

    def set_model(self, model):
        self.model = model

    def set_params(self, params):
        self.params = params


callbacks = types.SimpleNamespace(Callback=Callback)


class SGD:
    """The smallest optimizer with Keras' surface: config round trip, learning_rate variable,
    apply_gradients over (gradient, Variable) pairs."""

    def __init__(self, learning_rate=0.01):
        self.learning_rate = Variable(learning_rate)
        self.applied = 0

    def get_config(self):
        return {"learning_rate": float(self.learning_rate.value)}

    @classmethod
    def from_config(cls, config):
        return cls(**config)

    def variables(self):
        return []

    def apply_gradients(self, grads_and_vars):
        lr = float(self.learning_rate.value)
        for g, v in grads_and_vars:
            v.assign(v.value - lr * g)
        self.applied += 1


optimizers = types.SimpleNamespace(SGD=SGD)


class Model:
    def __init__(self, variables, optimizer):
        self.variables, self.optimizer = list(variables), optimizer
