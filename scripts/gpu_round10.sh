#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python scripts/mg_wingbox.py > gpurun_out/mg_wingbox.log 2>&1; echo "rc=$?"; tail -n 20 gpurun_out/mg_wingbox.log
