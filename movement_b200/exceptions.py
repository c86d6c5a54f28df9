"""Error types of the drop-in surface."""


class ConvergenceError(Exception):
    """Same name and role as ``firedrake.exceptions.ConvergenceError`` (movement/mover.py:116-132)."""
