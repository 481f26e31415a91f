"""Reference-compatible import paths: ``models.layers``, ``models.comm_ops``, ``models.model``."""
