"""TEST INFRASTRUCTURE - empty stand-in package so `import matplotlib.pyplot` in
the reference's SSCS_maker.py (line 65) succeeds; the PNG plot is out of scope."""
