"""Reference-compatible import paths: ``models.layers`` (reference models/layers.py:14-155), ``models.comm_ops``
(models/comm_ops.py:7-83), ``models.model`` (models/model.py:17-158)."""
