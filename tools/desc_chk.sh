#!/bin/bash
timeout 400 python -m pytest tests/test_gpu_extractor.py tests/test_gpu_wire.py tests/test_gpu_golden.py tests/test_gpu_ref.py -m gpu -x -q 2>&1 | tail -2 | cut -c1-300
bash tools/desc_ab.sh X=1
