#!/usr/bin/env bash
# The analysis scripts need pandas; matplotlib is optional (plots fall back to a PIL renderer).
set -euo pipefail
if python -c "import pandas, numpy, PIL" 2>/dev/null; then echo "analysis dependencies present"; else pip3 install --user pandas numpy pillow; fi
if python -c "import matplotlib" 2>/dev/null; then echo "matplotlib present"; else echo "matplotlib absent: plots will use the built-in PIL renderer"; fi
