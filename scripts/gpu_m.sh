#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/qs_ab_long.py > gpurun_out/m_qs_long.txt 2>&1; cat gpurun_out/m_qs_long.txt | tail -8
