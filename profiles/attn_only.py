"""Attention microbench only (through the C ABI): python profiles/attn_only.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from attn_microbench import attn
attn(16, 8, 1037)      # encoder, 8 views
attn(12, 8, 1036)      # frame, 8 views
attn(12, 1, 8289)      # global, 8 views
attn(12, 1, 66305)     # global, 64 views
