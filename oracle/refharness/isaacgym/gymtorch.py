"""Stub gymtorch: tensors are already torch tensors (zero-copy identity)."""
def wrap_tensor(t):
    return t
def unwrap_tensor(t):
    return t
