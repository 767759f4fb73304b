#!/bin/bash
set -u
O=gpurun_out/r2ac
mkdir -p $O
timeout 200 python profiles/scripts/with_lib.py ctrace profiles/scripts/chain_trace.py 360 0 60 > $O/trace_cfg5.txt 2>&1; echo "trace exit $?"; cat $O/trace_cfg5.txt | tail -50
timeout 200 python profiles/scripts/with_lib.py ctrace profiles/scripts/chain_trace.py 16 0 > $O/trace_cfg2.txt 2>&1; echo "trace exit $?"; cat $O/trace_cfg2.txt | tail -25
