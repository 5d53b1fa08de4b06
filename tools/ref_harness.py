"""Build-container-only helper: import the UNMODIFIED reference modules from /root/reference/Code
with matplotlib / skimage stubbed (they are absent here and only used for figures), so that
tools/make_golden.py can run the reference and record golden outputs.  Never used at test /
bench / product time (the GPU box has no /root/reference)."""
import os, sys
from unittest.mock import MagicMock

REF = "/root/reference"


def import_reference():
    for n in ["matplotlib", "matplotlib.pyplot", "mpl_toolkits", "mpl_toolkits.mplot3d", "skimage",
              "skimage.restoration"]:
        sys.modules.setdefault(n, MagicMock())
    plt = sys.modules["matplotlib.pyplot"]
    sys.modules["matplotlib"].pyplot = plt
    plt.subplots = lambda *a, **k: (MagicMock(), MagicMock())
    sys.path.insert(0, os.path.join(REF, "Code"))
    os.chdir(REF)
    import warnings
    warnings.filterwarnings("ignore")
    import shape_ftp, force_sensor, temperature_sensor
    return shape_ftp, force_sensor, temperature_sensor
