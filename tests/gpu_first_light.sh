#!/bin/bash
# first GPU contact: smoke + parity tests
set -x
nvidia-smi -L
python __graft_entry__.py smoke 2>&1 | tail -5
python -m pytest tests -m gpu -x -q -s 2>&1 | tail -40
