"""Drop-in for the reference's models/model_res_sigmoid.py (single environment, e.g. UR5 d=6):
`NN(device, dim, B)` and `Model(ModelPath, DataPath, dim, device)` with the same methods."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pntf.nn import NN_single as NN  # noqa: E402,F401
from pntf.model import Model_single as Model  # noqa: E402,F401
