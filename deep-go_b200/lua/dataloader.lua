-- dataloader.lua of deep-go, re-hosted on libdeepgo_b200.so.  Same globals as the reference
-- (STONE..TOTAL, SIZE, INPUT, input_dimensions; dataloader.lua:6-27) and the same functions
-- (flatten_data, transform, preprocess, empty_minibatch_io, load_minibatch, load_and_preprocess);
-- records are read by dg_t7_read_record and encoded on the GPU by dg_encode -- no torch needed.
local dg = require 'deepgo'
local ffi, lib = dg.ffi, dg.lib

STONE = 1
LIBERTIES = STONE + 3
LIBERTIES_AFTER = LIBERTIES + 4
KILL = LIBERTIES_AFTER + 7
AGE = KILL + 7
LADDER = AGE + 5
RANK = LADDER + 1
TOTAL = RANK + 9
SIZE = 19
input_dimensions = {TOTAL, SIZE, SIZE}
output_dimensions = {}
INPUT = {stone = 1, liberties = 2, liberties_after = 3, kills = 5, age = 7, ladder = 8, total = 9}

local RAW_BYTES = INPUT.total * SIZE * SIZE
local PLANE_FLOATS = TOTAL * SIZE * SIZE
local loader_ctx

local function context()
    loader_ctx = loader_ctx or dg.context(0)
    return loader_ctx
end

-- board summary (tables of 19x19 numbers) -> 9-plane byte record (dataloader.lua:30-39)
function flatten_data(data)
    local raw = ffi.new('uint8_t[?]', RAW_BYTES)
    local function put(plane, grid)
        for i = 1, SIZE do for j = 1, SIZE do raw[((plane - 1) * SIZE + i - 1) * SIZE + j - 1] = grid[i][j] % 256 end end
    end
    put(INPUT.stone, data.stones)
    put(INPUT.liberties, data.liberties)
    local la = data.liberties_after or data.liberties_after_move
    put(INPUT.liberties_after, la[1]); put(INPUT.liberties_after + 1, la[2])
    put(INPUT.kills, data.kills[1]); put(INPUT.kills + 1, data.kills[2])
    put(INPUT.age, data.age or data.ages)
    if data.ladders then put(INPUT.ladder, data.ladders[1]); put(INPUT.ladder + 1, data.ladders[2]) end
    return {input = raw, move = data.move, ranks = data.ranks, flat = true}
end

-- symmetry hook: the reference returns the grid unchanged (dataloader.lua:41-44); here the symmetry id travels
-- with the record and is applied inside the encode kernel (id 0 = reference behaviour)
function transform(grid)
    return grid
end

local function read_record(filename)
    local raw = ffi.new('uint8_t[?]', RAW_BYTES)
    local x, y, p, r = ffi.new('int[1]'), ffi.new('int[1]'), ffi.new('int[1]'), ffi.new('int[2]')
    dg.check(lib.dg_t7_read_record(filename, raw, x, y, p, r))
    return {input = raw, move = {x = x[0], y = y[0], player = p[0]}, ranks = {r[0], r[1]}, flat = true}
end

-- one record -> {input = float[37*19*19], output = 19*(x-1)+y} (dataloader.lua:50-92)
function preprocess(data, sym)
    local player = ffi.new('uint8_t[1]', data.move.player)
    local rank = ffi.new('uint8_t[1]', data.ranks[data.move.player])
    local s = ffi.new('uint8_t[1]', sym or 0)
    local out = ffi.new('float[?]', PLANE_FLOATS)
    dg.check(lib.dg_encode(context(), transform(data.input), player, rank, s, 1, out, dg.FMT_F32_NCHW))
    local xs, ys = ffi.new('uint8_t[1]', data.move.x), ffi.new('uint8_t[1]', data.move.y)
    local label = ffi.new('int32_t[1]')
    dg.check(lib.dg_encode_labels(xs, ys, s, 1, label))
    return {input = out, output = label[0] + 1}
end

function empty_minibatch_io(size)
    return ffi.new('float[?]', size * PLANE_FLOATS), ffi.new('double[?]', size)
end

-- read + encode a list of record files in ONE device call (dataloader.lua:113-121)
function load_minibatch(filenames)
    local n = #filenames
    local inputs, outputs = empty_minibatch_io(n)
    local raw = ffi.new('uint8_t[?]', n * RAW_BYTES)
    local player, rank = ffi.new('uint8_t[?]', n), ffi.new('uint8_t[?]', n)
    local xs, ys = ffi.new('uint8_t[?]', n), ffi.new('uint8_t[?]', n)
    for i = 1, n do
        local rec = read_record(filenames[i])
        ffi.copy(raw + (i - 1) * RAW_BYTES, rec.input, RAW_BYTES)
        player[i - 1], rank[i - 1] = rec.move.player, rec.ranks[rec.move.player]
        xs[i - 1], ys[i - 1] = rec.move.x, rec.move.y
    end
    if n > 0 then
        dg.check(lib.dg_encode(context(), raw, player, rank, nil, n, inputs, dg.FMT_F32_NCHW))
        local labels = ffi.new('int32_t[?]', n)
        dg.check(lib.dg_encode_labels(xs, ys, nil, n, labels))
        for i = 0, n - 1 do outputs[i] = labels[i] + 1 end
    end
    return {input = inputs, output = outputs, size = n, raw = raw, player = player, rank = rank, x = xs, y = ys}
end

function load_and_preprocess(filename)
    return preprocess(read_record(filename))
end
