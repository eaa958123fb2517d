#!/bin/bash
# the no-allocation entry point: its test alone in a fresh process, then the whole GPU suite (it is collected last)
out=gpurun_out/noalloc; mkdir -p $out
( time python -m pytest tests/test_zz_no_alloc.py -m gpu -q ) 2>&1 | cut -c1-400 | tail -40 > $out/alone.log; cat $out/alone.log
( time python -m pytest tests -m gpu -x -q --durations=6 ) 2>&1 | grep -E "^E  |FAILED|passed|failed|Error|^real|s call" | cut -c1-300 | head -60 > $out/suite.log; cat $out/suite.log
