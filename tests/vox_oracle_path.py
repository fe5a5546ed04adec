"""Puts oracle/ on sys.path for tests that import vox_oracle (test infrastructure only)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
