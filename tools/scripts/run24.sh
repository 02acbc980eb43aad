#!/bin/bash
cd /root/repo
CPN_MLP_TS=1 CPN_MLP_DBG=512 timeout 120 python tools/run_op.py mlp --M 200704 --C 96 --iters 1 2>&1 | tail -32
CPN_MLP_TS=1 CPN_MLP_DBG=256 timeout 120 python tools/run_op.py mlp --M 50176 --C 192 --iters 1 2>&1 | tail -36
