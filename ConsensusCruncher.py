#!/usr/bin/env python
"""Drop-in entry point: `ConsensusCruncher.py [-c CONFIG] consensus -i BAM -o OUT ...` on the B200 path
(see consensuscruncher_b200/cli.py; flags of the reference's ConsensusCruncher.py:433-443)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from consensuscruncher_b200.cli import main  # noqa: E402

if __name__ == "__main__":
    sys.exit(main())
