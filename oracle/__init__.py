"""TEST INFRASTRUCTURE.  CPU restatement of the reference's consensus path and
the harness that runs the unmodified reference.  Only tests/, smoke() and
bench.py's cpu_baseline / --impl reference legs may import this package."""
