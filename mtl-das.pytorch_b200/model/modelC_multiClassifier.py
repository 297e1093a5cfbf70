"""Same import path as the reference (`from model.modelC_multiClassifier import Multi_Classifier`, utils.py:20)."""
from mtldas_b200.nets_c import Multi_Classifier  # noqa: F401
