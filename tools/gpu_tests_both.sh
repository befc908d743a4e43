#!/bin/bash
# the GPU parity suite twice: as built, and with three frames in flight forced on for every context (RT_PIPELINE=1)
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
RT_PIPELINE=1 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
