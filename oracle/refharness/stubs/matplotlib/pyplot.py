"""Empty matplotlib.pyplot stand-in (test infrastructure only)."""
