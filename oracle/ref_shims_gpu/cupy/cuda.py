"""cupy.cuda stand-in (tests only): Device / runtime / Stream over the gt4py_b200 C ABI."""
from gt4py_b200 import _lib


class Device:
    def __init__(self, device=None):
        if _lib.device_count() < 1:
            raise RuntimeError("cupy stand-in: no CUDA device")
        self.id = 0 if device is None else int(device)

    def synchronize(self):
        _lib.synchronize()

    def use(self):
        _lib.set_device(self.id)
        return self

    @property
    def compute_capability(self):
        major, minor = _lib.device_props()["cc"]
        return f"{major}{minor}"

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


class _Runtime:
    is_hip = False

    @staticmethod
    def getDeviceCount():
        return _lib.device_count()

    @staticmethod
    def deviceSynchronize():
        _lib.synchronize()


runtime = _Runtime()


class Stream:
    null = None

    def __init__(self, null=False, non_blocking=False):
        self._s = None if null else _lib.Stream()
        self.ptr = 0 if null else int(self._s.handle.value)

    def synchronize(self):
        if self._s is None:
            _lib.synchronize()
        else:
            self._s.synchronize()


def get_current_stream():
    return Stream(null=True)
