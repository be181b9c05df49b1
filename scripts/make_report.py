#!/usr/bin/env python3
"""CLI shim at the reference path scripts/make_report.py -> b200bench.analysis.make_report."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200bench.analysis.make_report import main

if __name__ == "__main__":
    main()
