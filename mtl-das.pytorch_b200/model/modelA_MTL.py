"""Same import path as the reference (`from model.modelA_MTL import MTL_Net`, utils.py:18)."""
from mtldas_b200.nets import MTL_Net  # noqa: F401
