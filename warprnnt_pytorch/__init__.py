"""Drop-in for the import at /root/reference rnnt/model.py:6
(``from warprnnt_pytorch import RNNTLoss``): with this repo's root on ``sys.path`` the
reference's ``Transducer`` constructs the B200-native loss without any code change."""
from rnnt_transducer_b200 import RNNTLoss, rnnt_loss  # noqa: F401

__all__ = ["RNNTLoss", "rnnt_loss"]
