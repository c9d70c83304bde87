"""Drop-in for the reference's models/model_res_sigmoid_multi.py (multi environment, Gibson/C3D d=3):
`NN(device, dim)` and `Model(ModelPath, DataPath, dim, length, device)` with the same methods."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pntf.nn import NN_multi as NN  # noqa: E402,F401
from pntf.model import Model_multi as Model  # noqa: E402,F401
