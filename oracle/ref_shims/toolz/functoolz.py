import functools
import inspect


def identity(x):
    return x


def complement(func):
    return lambda *a, **k: not func(*a, **k)


class curry:
    def __init__(self, func, *args, **kwargs):
        self.func = func
        self.args = args
        self.keywords = kwargs
        functools.update_wrapper(self, func)

    def __call__(self, *args, **kwargs):
        args = self.args + args
        kw = {**self.keywords, **kwargs}
        try:
            sig = inspect.signature(self.func)
            sig.bind(*args, **kw)
        except TypeError:
            return curry(self.func, *args, **kw)
        except ValueError:
            pass
        return self.func(*args, **kw)
