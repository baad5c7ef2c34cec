class DistributedDataParallelKwargs:
    def __init__(self, *a, **k):
        pass
