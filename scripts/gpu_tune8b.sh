#!/bin/bash
for mt in 1 2 4; do for nw in 8 16; do RT_GEMV_MT=$mt RT_GEMV_NW=$nw python scripts/gemv_tune8b.py; done; done
python scripts/gemv_tune8b.py
