"""Build tools/tune_adam.cu once per cache-policy variant (tools/hint_<X>.bin) for the hint sweep."""
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
NA, EF, EL = ".L1::no_allocate", ".L1::no_allocate.L2::evict_first", ".L1::no_allocate.L2::evict_last"
VARIANTS = {  # name: (ld narrow, st narrow, ld 256-bit, st 256-bit)
    "A": (NA, NA, NA, NA),        # product default
    "B": (NA, NA, EF, EF),
    "C": (NA, NA, NA, EF),
    "D": ("", "", "", ""),       # no hints at all
    "E": (NA, NA, EF, NA),
    "F": (".cs", ".cs", ".cs", ".cs"),
    "G": (NA, NA, EL, EF),
}
for name, (ld, st, ld256, st256) in VARIANTS.items():
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-fmad=false",
           f'-DMT_LD_HINT="{ld}"', f'-DMT_ST_HINT="{st}"', f'-DMT_LD_HINT256="{ld256}"', f'-DMT_ST_HINT256="{st256}"',
           "-o", str(HERE / f"hint_{name}.bin"), str(HERE / "tune_adam.cu")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    print(name, "ok" if r.returncode == 0 else "FAILED\n" + r.stderr[-400:])
    if r.returncode:
        sys.exit(1)
