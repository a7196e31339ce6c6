"""python -m tahoe_b200.peaks -> one JSON line: the L2 / shared-memory peaks of the current GPU."""
import json

from . import api

if __name__ == "__main__":
    print(json.dumps(api.measure_peaks()))
