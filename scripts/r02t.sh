#!/bin/bash
python scripts/small_batch_probe.py 2>&1 | tail -8
