"""Pack the reference's fixture corpus into tests/golden/positions.npz.

Run HERE (the container that has /root/reference); the GPU box never sees the reference.
    python -m oracle.make_golden
Reads every record listed by data/{train,validation,test}_game_counts.txt
(data.lua:53-80 index format) with oracle.t7 and stores, in file order:
  raw[N,9,19,19] u8, x,y,player[N] (1-based as in the records), ranks[N,2], game[N] (index
  into games), move[N] (1-based file name), split[N] (0 train, 1 validation, 2 test), games (names).
These are DATA golden vectors (outputs of the reference's makedata.lua), not source.
"""
import os
import sys
import numpy as np
from . import t7

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "positions.npz")


def main():
    raws, xs, ys, ps, rks, gi, mi, sp, games = [], [], [], [], [], [], [], [], []
    for s, split in enumerate(("train", "validation", "test")):
        with open(os.path.join(REF, "data", split + "_game_counts.txt")) as f:
            for line in f:
                name, cnt = line.rstrip("\n").split("\t")
                cnt = int(cnt)
                if cnt <= 0:          # data.lua:74 drops empty games
                    continue
                g = len(games)
                games.append(name)
                for m in range(1, cnt + 1):
                    raw, x, y, p, rk = t7.load_record(os.path.join(REF, name, str(m)))
                    raws.append(raw); xs.append(x); ys.append(y); ps.append(p); rks.append(rk)
                    gi.append(g); mi.append(m); sp.append(s)
    np.savez_compressed(
        OUT, raw=np.stack(raws), x=np.array(xs, np.uint8), y=np.array(ys, np.uint8),
        player=np.array(ps, np.uint8), ranks=np.array(rks, np.uint8), game=np.array(gi, np.int32),
        move=np.array(mi, np.int32), split=np.array(sp, np.uint8), games=np.array(games))
    print("wrote", OUT, len(raws), "records", os.path.getsize(OUT), "bytes")
    # a handful of records in their original Torch7 wire format (3,598 B each) for the record-reader tests,
    # laid out like the reference's data/ tree with its index file (data.lua:53-80)
    import shutil
    rec_root = os.path.join(os.path.dirname(OUT), "records")
    game = "data/test/1993/2000-03-24b.sgf"
    os.makedirs(os.path.join(rec_root, game), exist_ok=True)
    for m in (1, 2, 3, 60, 125):
        shutil.copyfile(os.path.join(REF, game, str(m)), os.path.join(rec_root, game, str(m)))
    with open(os.path.join(rec_root, "data", "test_game_counts.txt"), "w") as f:
        f.write("%s\t%8d\n" % (game, 3))          # moves 1..3 are present consecutively
        f.write("data/test/1993/empty.sgf\t%8d\n" % 0)


if __name__ == "__main__":
    sys.exit(main())
