#!/bin/bash
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r26_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/r26_tests.log
