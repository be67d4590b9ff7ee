#!/usr/bin/env python
"""Drop-in name of the reference's script (README:12): same flags, same printed lines, B200 engine."""
import runpy
import os

runpy.run_path(os.path.join(os.path.dirname(os.path.abspath(__file__)), "train.py"), run_name="__main__")
