import sys, numpy as np
sys.path.insert(0, '.')
import hoomd_b200 as hb
from dbg_modes import mk
