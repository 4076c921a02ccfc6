class DDPStrategy:
    def __init__(self, *a, **k):
        self.kwargs = k
