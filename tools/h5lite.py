"""Moved to igm_b200/h5lite.py; kept so that tests/golden/make_golden.py's import keeps working."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from igm_b200.h5lite import *          # noqa: F401,F403,E402
from igm_b200.h5lite import File       # noqa: F401,E402
