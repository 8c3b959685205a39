"""Import shim for deepdiff (placeholders; never used on the numeric path)."""


class _Diff:
    class DeepDiff(dict):
        def __init__(self, old=None, new=None, **kwargs):
            super().__init__({} if old == new else {"values_changed": (old, new)})


class _Hash:
    class DeepHash(dict):
        def __init__(self, obj, **kwargs):
            super().__init__()
            self._obj = obj

        def __getitem__(self, obj):
            return hash(repr(obj))


diff = _Diff
deephash = _Hash
DeepDiff = _Diff.DeepDiff
DeepHash = _Hash.DeepHash
