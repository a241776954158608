-- makedata.lua of deep-go, re-hosted on libdeepgo_b200.so: SGF text -> records.  The string work (all_moves, handicaps,
-- get_ranks, makedata.lua:24-120) is dg_sgf_parse; the replay (all_boards :156-186) plays every game's moves with
-- dg_update_boards and summarises all positions of a game in ONE dg_summarize_boards call.  Authored for LuaJIT; the
-- executable mirror is deep-go_b200/host/makedata.py (many games in lockstep).
local ffi = require 'ffi'
local deepgo = require 'deepgo'
local lib, check = deepgo.lib, deepgo.check

local MAX_MOVES, MAX_HANDICAPS = 1024, 64

function parse_sgf(s)
    local ranks, has = ffi.new("int[2]"), ffi.new("int[1]")
    local h = {ffi.new("uint8_t[?]", MAX_HANDICAPS), ffi.new("uint8_t[?]", MAX_HANDICAPS), ffi.new("uint8_t[?]", MAX_HANDICAPS)}
    local m = {ffi.new("uint8_t[?]", MAX_MOVES), ffi.new("uint8_t[?]", MAX_MOVES), ffi.new("uint8_t[?]", MAX_MOVES)}
    local nh, nm = ffi.new("int[1]"), ffi.new("int[1]")
    check(lib.dg_sgf_parse(s, #s, ranks, has, h[1], h[2], h[3], MAX_HANDICAPS, nh, m[1], m[2], m[3], MAX_MOVES, nm))
    return {ranks = has[0] ~= 0 and {ranks[0], ranks[1]} or nil, handicaps = h, n_handicaps = nh[0], moves = m, n_moves = nm[0]}
end

-- transcribe_from_to (makedata.lua:537-559) for one game: returns n, raw uint8[n][9][19][19], and the parsed game
function transcribe_game(ctx, s)
    local g = parse_sgf(s)
    if g.ranks == nil or g.n_moves == 0 then return 0, nil, g end
    local n = g.n_moves
    local stones, age = ffi.new("uint8_t[361]"), ffi.new("uint8_t[361]")
    for i = 0, g.n_handicaps - 1 do
        check(lib.dg_update_boards(ctx, stones, age, g.handicaps[1] + i, g.handicaps[2] + i, g.handicaps[3] + i, 1))
    end
    local all_stones, all_age = ffi.new("uint8_t[?]", n * 361), ffi.new("uint8_t[?]", n * 361)
    for i = 0, n - 1 do
        ffi.copy(all_stones + i * 361, stones, 361)            -- the board BEFORE the move
        ffi.copy(all_age + i * 361, age, 361)
        check(lib.dg_update_boards(ctx, stones, age, g.moves[1] + i, g.moves[2] + i, g.moves[3] + i, 1))
    end
    local summ = ffi.new("uint8_t[?]", n * 7 * 361)
    check(lib.dg_summarize_boards(ctx, all_stones, summ, n))
    local raw = ffi.new("uint8_t[?]", n * 9 * 361)             -- flatten_data plane order (dataloader.lua:20-39)
    for i = 0, n - 1 do
        ffi.copy(raw + (i * 9 + 0) * 361, all_stones + i * 361, 361)
        ffi.copy(raw + (i * 9 + 1) * 361, summ + (i * 7 + 0) * 361, 5 * 361)
        ffi.copy(raw + (i * 9 + 6) * 361, all_age + i * 361, 361)
        ffi.copy(raw + (i * 9 + 7) * 361, summ + (i * 7 + 5) * 361, 2 * 361)
    end
    return n, raw, g
end
