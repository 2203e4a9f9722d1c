timeout 600 python tools/cascade_probe.py 2>&1 | tail -12
