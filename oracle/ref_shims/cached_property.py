from functools import cached_property  # import shim
