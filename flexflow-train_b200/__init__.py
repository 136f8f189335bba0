"""flexflow-train_b200: B200-native (sm_100a) per-operator training kernels behind
FlexFlow-train's lib/kernels API. The directory name carries a hyphen (it is the name the
project was given), so import it through `ffk_import.py` at the repo root, which registers
it as the module `flexflow_train_b200`.
"""
from ._lib import FFKError, lib, parse_headers, LIB_PATH  # noqa: F401
