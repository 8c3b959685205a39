def debug(*args, **kwargs):  # import shim
    return None
