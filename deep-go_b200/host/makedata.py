"""Host mirror of makedata.lua's corpus build: SGF text -> the 9-plane records.

    parse(sgf)                      all_moves / handicaps / get_ranks (makedata.lua:24-120)  -> dg_sgf_parse (host C++)
    all_boards(games, boards)       makedata.lua:156-186: summarise the board BEFORE each move, record the age BEFORE
                                    the move, then play it -- for MANY games in lockstep: step t plays move t of every game
                                    that still has one, in one dg_update_boards call, and all positions of all games are
                                    summarised in one dg_summarize_boards call at the end
    transcribe(sgf_texts, boards)   transcribe_from_to (:537-559) for a list of games -> arrays of records
    transcribe_to_corpus(...)       ... written as a packed corpus (host/data.py) instead of one Torch7 file per position

`boards` is anything with update_boards(stones, age, x, y, player) and summarize_boards(stones): an engine.Context (GPU).
"""
import numpy as np

from deepgo_b200 import check, ffi, lib

MAX_MOVES = 1024
MAX_HANDICAPS = 64


def parse(sgf):
    """-> dict(ranks=(r1, r2) or None, handicaps=[(x, y, player)], moves=[(x, y, player)]), 1-based, player 1 = black."""
    data = sgf if isinstance(sgf, bytes) else sgf.encode("latin-1")
    ranks, has = ffi.new("int[2]"), ffi.new("int *")
    hx, hy, hp = (ffi.new("uint8_t[]", MAX_HANDICAPS) for _ in range(3))
    mx, my, mp = (ffi.new("uint8_t[]", MAX_MOVES) for _ in range(3))
    nh, nm = ffi.new("int *"), ffi.new("int *")
    check(lib().dg_sgf_parse(data, len(data), ranks, has, hx, hy, hp, MAX_HANDICAPS, nh, mx, my, mp, MAX_MOVES, nm))
    return {"ranks": (ranks[0], ranks[1]) if has[0] else None,
            "handicaps": [(hx[i], hy[i], hp[i]) for i in range(nh[0])],
            "moves": [(mx[i], my[i], mp[i]) for i in range(nm[0])]}


def _play(boards, stones, age, games, which, step):
    """play move `step` of list `which` on every game that has one (one lockstep update_boards call)"""
    live = [g for g in range(len(games)) if step < len(games[g][which])]
    if not live:
        return False
    mv = np.array([games[g][which][step] for g in live], np.uint8)
    s, a = np.ascontiguousarray(stones[live]), np.ascontiguousarray(age[live])
    boards.update_boards(s, a, mv[:, 0], mv[:, 1], mv[:, 2])
    stones[live], age[live] = s, a
    return True


def all_boards(games, boards):
    """games: list of parse() results -> per game the (stones, age) BEFORE each of its moves, [n_moves, 19, 19] uint8 each."""
    n = len(games)
    stones, age = np.zeros((n, 19, 19), np.uint8), np.zeros((n, 19, 19), np.uint8)
    step = 0
    while _play(boards, stones, age, games, "handicaps", step):
        step += 1
    snaps = [([], []) for _ in range(n)]
    step = 0
    while True:
        for g in range(n):
            if step < len(games[g]["moves"]):
                snaps[g][0].append(stones[g].copy()); snaps[g][1].append(age[g].copy())
        if not _play(boards, stones, age, games, "moves", step):
            break
        step += 1
    empty = np.zeros((0, 19, 19), np.uint8)
    return [(np.stack(s) if s else empty, np.stack(a) if a else empty) for s, a in snaps]


def transcribe(sgf_texts, boards):
    """-> dict(raw [N,9,19,19], x, y, player [N], ranks [N,2], game [N]): the records transcribe_from_to would save, in game
    order then move order.  Games without two dan ranks give no records (makedata.lua:548-556)."""
    games = [parse(s) for s in sgf_texts]
    keep = [i for i, g in enumerate(games) if g["ranks"] is not None]
    snaps = all_boards([games[i] for i in keep], boards)
    stones = np.concatenate([s for s, _ in snaps]) if snaps else np.zeros((0, 19, 19), np.uint8)
    age = np.concatenate([a for _, a in snaps]) if snaps else np.zeros((0, 19, 19), np.uint8)
    n = len(stones)
    raw = np.zeros((n, 9, 19, 19), np.uint8)
    if n:
        summ = boards.summarize_boards(stones)           # liberties, liberties_after[2], kills[2], ladders[2]
        raw[:, 0] = stones
        raw[:, 1:6] = summ[:, 0:5]
        raw[:, 6] = age
        raw[:, 7:9] = summ[:, 5:7]                       # flatten_data plane order (dataloader.lua:20-39)
    mv = np.array([m for i in keep for m in games[i]["moves"]], np.uint8).reshape(-1, 3)
    rk = np.array([games[i]["ranks"] for i in keep for _ in games[i]["moves"]], np.uint8).reshape(-1, 2)
    gm = np.array([i for i in keep for _ in games[i]["moves"]], np.int64)
    return {"raw": raw, "x": mv[:, 0], "y": mv[:, 1], "player": mv[:, 2], "ranks": rk, "game": gm}


def transcribe_to_corpus(sgf_texts, boards, out_path):
    from . import data
    r = transcribe(sgf_texts, boards)
    data.write_corpus_from_arrays(out_path, r["raw"], r["x"], r["y"], r["player"], r["ranks"])
    return len(r["raw"])
