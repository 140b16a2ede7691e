import os, sys
sys.path.insert(0, os.getcwd())
sys.path.insert(0, os.path.join(os.getcwd(), "profiles", "ubench"))
import kernel_bench as kb
kb.bench_attn("long self (12 x 64) 8x4000", 8, 4000, 12, 64, False, True)
kb.bench_attn("short self (12 x 64) 256x250", 256, 250, 12, 64, False, True)
