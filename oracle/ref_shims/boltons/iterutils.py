"""Import shim for boltons.iterutils."""


def is_collection(obj):
    if isinstance(obj, (str, bytes)):
        return False
    try:
        iter(obj)
    except TypeError:
        return False
    return True


def flatten_iter(iterable):
    for item in iterable:
        if is_collection(item):
            yield from flatten_iter(item)
        else:
            yield item


def flatten(iterable):
    return list(flatten_iter(iterable))
