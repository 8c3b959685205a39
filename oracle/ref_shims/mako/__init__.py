"""Import shim for mako."""
