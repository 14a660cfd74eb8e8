#!/bin/bash
python scripts/merge_local_probe.py 2>&1 | tail -14
