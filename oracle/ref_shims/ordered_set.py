class OrderedSet(dict):  # import shim
    def __init__(self, it=()):
        super().__init__((k, None) for k in it)

    def add(self, k):
        self[k] = None
