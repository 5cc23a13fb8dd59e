#!/bin/bash
# builds libtutor_b200.so in-tree; prints ptxas summaries of kernels matching $1 (default: errors only)
cd "$(dirname "$0")/.." && python -c "
import importlib,sys
sys.path.insert(0,'.')
b=importlib.import_module('distributed-lms-raft-llm_b200.build')
print(b.build(verbose=True))" 2>&1 | grep -B2 -A6 "${1:-rror}\|rror\|setmaxnreg" | head -${2:-60}
