"""Minimal stand-in for the `keras` package so that the reference's setups (`import keras.backend`,
e.g. code/setups/fixpoint-density.py:10) import unchanged.  Only `keras.backend.clear_session` exists:
there is no TensorFlow graph to clear -- nets live in the engine's device arena."""
from . import backend  # noqa: F401
