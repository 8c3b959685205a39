"""Import shim (test infrastructure only): lets the read-only reference import in this container."""
