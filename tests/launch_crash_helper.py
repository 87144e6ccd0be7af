"""Stand-in training script for the launcher's crash-recovery test: rank 1 dies on the first attempt (the one started
without ``--resume``); on the restart every rank sees ``--resume latest`` and finishes.  The attempt is told apart by the
argument, not by the marker file, so a rank that starts late cannot mistake the first attempt for the second."""
import os
import sys
import time
from pathlib import Path


def main() -> None:
    marker = Path(sys.argv[sys.argv.index("--marker") + 1])
    rank = int(os.environ.get("RANK", "0"))
    resumed = "--resume" in sys.argv and sys.argv[sys.argv.index("--resume") + 1] == "latest"
    if not resumed:
        if rank == 1:
            marker.write_text("crashed once")
            sys.exit(7)
        time.sleep(60)          # the surviving rank would hang in a collective: the launcher must take it down
        sys.exit(0)
    assert marker.exists(), "a restart without a first attempt"
    (marker.parent / f"done_{rank}").write_text(os.environ["MASTER_PORT"])


if __name__ == "__main__":
    main()
