"""ORACLE (test infrastructure only): plain-Python restatement of makedata.lua's SGF string handling.
Parity unpinned: the reference ships no SGF file and Lua cannot run here; every function follows the cited lines 1:1,
including the behaviour of its custom string:split (makedata.lua:14-22)."""

CHARS = "abcdefghijklmnopqrs"                     # makedata.lua:7-12


def split(s, sep):
    """string:split(pat) for a literal separator: all segments, empty ones included, remainder last (:14-22)."""
    out, st = [], 0
    while True:
        i = s.find(sep, st)
        if i < 0:
            out.append(s[st:])
            return out
        out.append(s[st:i])
        st = i + len(sep)


def split_sgf(s):
    """(name, value) tokens (:40-57): lines at CR LF, pieces at ';', exactly two '['-parts, the second ending in ']'."""
    result = []
    for line in split(s, "\r\n"):
        for piece in split(line, ";"):
            sub = split(piece, "[")
            if len(sub) == 2 and sub[1][-1:] == "]":
                result.append((sub[0], sub[1][:-1]))
    return result


def to_move(s):
    """:59-66 -> (x, y) with None for characters outside a..s, or None if the value is not two characters"""
    if len(s) != 2:
        return None
    f = lambda c: CHARS.index(c) + 1 if c in CHARS else None
    return f(s[0]), f(s[1])


def all_moves(s):
    """:69-90 -> [(x, y, player)]; a move whose x is nil is skipped; y nil is kept (and breaks the reference later)"""
    out = []
    for name, value in split_sgf(s):
        if name in ("B", "W"):
            mv = to_move(value)
            if mv is not None and mv[0] is not None:
                out.append((mv[0], mv[1], 1 if name == "B" else 2))
    return out


def handicaps(s):
    """:24-38 -> [(x, y, player)] from LINES that start with AB / AW"""
    out = []
    for line in split(s, "\r\n"):
        if line[:2] in ("AB", "AW"):
            player = 2 if line[:2] == "AW" else 1
            for pos in split(line[3:-1] if len(line) >= 2 else "", "]["):
                mv = to_move(pos)
                out.append((mv[0] if mv else None, mv[1] if mv else None, player))
    return out


def to_rank(s):
    """:92-100: '<number>d' -> number, else None (Lua tonumber)"""
    if s[-1:] != "d":
        return None
    t = s[:-1].strip(" \t\n\r\f\v")
    try:
        return float(int(t, 16)) if t.lower().startswith(("0x", "-0x")) else float(t)
    except ValueError:
        return None


def get_ranks(s):
    """:102-120: the last BR / WR win; both must be dan ranks"""
    r = [None, None]
    for name, value in split_sgf(s):
        if name == "BR":
            r[0] = to_rank(value)
        elif name == "WR":
            r[1] = to_rank(value)
    return tuple(r) if r[0] is not None and r[1] is not None else None
