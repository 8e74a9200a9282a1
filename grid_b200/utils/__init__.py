"""Host-side mirrors of the reference's step modules (same module and function names as grid/utils/)."""
