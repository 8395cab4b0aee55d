#!/usr/bin/env python
"""`singleton_correction.py` with the reference's arguments, on the GPU path (see stages.py)."""
import os
import sys

if __package__ in (None, ""):
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from consensuscruncher_b200.stages import sc_main as main  # noqa: E402

if __name__ == "__main__":
    sys.exit(main())
