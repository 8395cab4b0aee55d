#!/usr/bin/env python3
"""`extract_barcodes.py --read1 R1 --read2 R2 --outfile PREFIX [--bpattern NNT | --blist FILE]`:
the reference's barcode extraction command (ConsensusCruncher/extract_barcodes.py:140-160)."""
import sys

from consensuscruncher_b200.extract import main

if __name__ == "__main__":
    sys.exit(main())
