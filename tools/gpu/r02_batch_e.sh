#!/bin/bash
# Round-2 GPU batch E: the allocation-free entry point alone in a fresh process, then the whole GPU suite (it runs last).
set -x
python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -4
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
