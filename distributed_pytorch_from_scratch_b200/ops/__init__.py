"""Python bindings and autograd wrappers for the hand-written sm_100a kernels."""
