import sys

from . import build

print(build(force="--force" in sys.argv, verbose=True))
