#!/bin/bash
set -u
O=gpurun_out/r02h
mkdir -p $O
python tools/research/latency_phases.py > $O/latency_phases.txt 2>&1; cat $O/latency_phases.txt
