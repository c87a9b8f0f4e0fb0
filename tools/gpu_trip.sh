#!/bin/bash
bash tools/tune_q1.sh q1
bash tools/ncu_q1_variants.sh
