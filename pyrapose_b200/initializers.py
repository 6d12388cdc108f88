"""Initialisers — mirrors PyraPose/initializers.py (PriorProbability, :23-39) and the Keras
initialisers the reference's model builders ask for (retinanet.py:35,42,78,85,106: normal(0, 0.01);
Keras Conv2D default glorot_uniform for the PFPN convs; keras-resnet convs are overwritten by the
ImageNet checkpoint in the reference, random here)."""
import math

import numpy as np


class PriorProbability(object):
    """Apply a prior probability to the weights: bias = -log((1 - p) / p)."""

    def __init__(self, probability=0.01):
        self.probability = probability

    def get_config(self):
        return {'probability': self.probability}

    def __call__(self, shape, dtype=None):
        return np.ones(shape, dtype=dtype or np.float32) * -math.log((1 - self.probability) / self.probability)


def normal(rng, shape, mean=0.0, stddev=0.01):
    return rng.normal(mean, stddev, size=shape).astype(np.float32)


def glorot_uniform(rng, shape):
    """Keras glorot_uniform for an HWIO kernel: limit = sqrt(6 / (fan_in + fan_out))."""
    kh, kw, cin, cout = shape
    limit = math.sqrt(6.0 / (kh * kw * cin + kh * kw * cout))
    return rng.uniform(-limit, limit, size=shape).astype(np.float32)


def he_normal(rng, shape):
    kh, kw, cin, _ = shape
    return rng.normal(0.0, math.sqrt(2.0 / (kh * kw * cin)), size=shape).astype(np.float32)
