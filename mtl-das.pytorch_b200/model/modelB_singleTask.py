"""Same import path as the reference (`from model.modelB_singleTask import Single_Task_Net`, utils.py:19)."""
from mtldas_b200.nets import Single_Task_Net  # noqa: F401
