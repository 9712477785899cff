python tools/dev_step_scaling.py 1920x1200 2>/dev/null
python tools/dev_step_scaling.py 752x480 2>/dev/null
