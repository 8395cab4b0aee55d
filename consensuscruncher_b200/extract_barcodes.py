#!/usr/bin/env python
"""`extract_barcodes.py --read1 R1 --read2 R2 --outfile PREFIX [--bpattern NNT | --blist FILE]` with the
reference's arguments (ConsensusCruncher/extract_barcodes.py:140-160); see extract.py."""
import os
import sys

if __package__ in (None, ""):
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from consensuscruncher_b200.extract import main  # noqa: E402

if __name__ == "__main__":
    sys.exit(main())
