import sys

from . import build


def main() -> None:
    print(build(force="--force" in sys.argv, verbose=True))


if __name__ == "__main__":
    main()
