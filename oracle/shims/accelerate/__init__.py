class Accelerator:  # inert stub (training only)
    def __init__(self, *a, **k):
        raise NotImplementedError
from . import utils  # noqa: E402,F401
