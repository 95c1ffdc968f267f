"""ctypes loader for the native library (placeholder until the first build)."""
import os

_LIB = None
LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "lib", "libdtf_b200.so")


def available() -> bool:
    return os.path.isfile(LIB_PATH)
