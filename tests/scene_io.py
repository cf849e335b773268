"""TEST INFRASTRUCTURE: build the reference's hg::Scene (through oracle/_ref) from a FlatScene, and run reference frames."""
import ctypes as C
import os

import numpy as np

from tests.oracle_lib import P


from harfang3d_b200.flatscene import SceneHandle as RefScene  # noqa: E402,F401  (one class drives the reference build and the facade)
