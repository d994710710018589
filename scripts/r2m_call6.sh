#!/bin/bash
mkdir -p gpurun_out
AB_ARGS="" bash scripts/ab.sh 2 - fakevec1 fakevec2 2>&1 | tee gpurun_out/m6_fakevec.txt
