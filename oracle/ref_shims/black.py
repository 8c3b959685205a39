"""Import shim for black: identity formatter."""


class _TV:
    def __getitem__(self, item):
        return item

    def __getattr__(self, item):
        return item


TargetVersion = _TV()


class FileMode:
    def __init__(self, **kwargs):
        self.kwargs = kwargs


Mode = FileMode


def format_str(src, mode=None):
    return src
