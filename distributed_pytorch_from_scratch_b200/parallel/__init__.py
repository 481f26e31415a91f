"""Tensor-parallel runtime: process groups, communication autograd ops, symmetric-memory communicator."""
