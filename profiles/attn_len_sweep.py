"""Global-attention lengths (2, 4, 8, 16 views): QT=1 (two CTAs/SM) vs QT=2 kernel, via MAB_ATT_SHORT_MAX."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from attn_microbench import attn
for n in (2, 4, 8, 16):
    attn(12, 1, n * 1036 + 1)
