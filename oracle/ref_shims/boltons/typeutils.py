"""Import shim for boltons.typeutils (no stencil arithmetic here)."""


class classproperty:
    def __init__(self, fn):
        self.fn = fn

    def __get__(self, obj, owner):
        return self.fn(owner)


def make_sentinel(name="_MISSING", var_name=None):
    class Sentinel:
        def __init__(self):
            self.name = name
            self.var_name = var_name

        def __repr__(self):
            return self.var_name or f"{type(self).__name__}({self.name!r})"

        def __reduce__(self):
            return self.var_name if self.var_name else super().__reduce__()

        def __bool__(self):
            return False

        def __copy__(self):
            return self

        def __deepcopy__(self, memo):
            return self

    Sentinel.__name__ = name
    return Sentinel()
