"""Kernel bindings (``native``) and autograd-visible wrappers (``functional``)."""
from . import native  # noqa: F401
