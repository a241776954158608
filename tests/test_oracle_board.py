"""Pins oracle/board.c on the reference's own fixture corpus (all 4,398 records)."""
import numpy as np

from oracle import board


def test_summarize_reproduces_every_fixture(golden):
    raw = golden["raw"]
    out = board.summarize(raw[:, 0])
    rec = raw[:, [1, 2, 3, 4, 5, 7, 8]]          # liberties, liberties_after x2, kills x2, ladders x2
    assert out.shape == rec.shape
    assert np.array_equal(out, rec)


def test_update_board_replays_every_successor(golden):
    raw, game, move = golden["raw"], golden["game"], golden["move"]
    pairs = 0
    for i in range(len(raw) - 1):
        if game[i + 1] != game[i] or move[i + 1] != move[i] + 1:
            continue
        s, a = raw[i, 0].copy(), raw[i, 6].copy()
        rc = board.update_board(s, a, int(golden["x"][i]) - 1, int(golden["y"][i]) - 1, int(golden["player"][i]))
        assert rc == 0
        assert np.array_equal(s, raw[i + 1, 0]) and np.array_equal(a, raw[i + 1, 6])
        pairs += 1
    assert pairs == 4376


def test_update_board_rejects_occupied():
    s = np.zeros((19, 19), np.uint8); a = np.zeros((19, 19), np.uint8)
    s[3, 3] = 1
    assert board.update_board(s, a, 3, 3, 2) == -1


def test_capture_and_suicide_rules():
    # white stone at corner with one liberty: black captures it (kills plane), suicide is 0 liberties-after
    s = np.zeros((19, 19), np.uint8)
    s[0, 0] = 2; s[0, 1] = 1
    out = board.summarize(s[None])[0]
    assert out[3, 1, 0] == 1            # kills[p=1] at (1,0)
    assert out[0, 0, 0] == 1            # liberties of the white stone
    s2 = np.zeros((19, 19), np.uint8)
    s2[0, 1] = 1; s2[1, 0] = 1
    out2 = board.summarize(s2[None])[0]
    assert out2[2, 0, 0] == 0           # white playing (0,0) is suicide: liberties_after[p=2] == 0
    assert out2[1, 0, 0] == 3           # black there connects both stones: liberties (0,2),(1,1),(2,0)
