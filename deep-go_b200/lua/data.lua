-- data.lua of deep-go, re-hosted: Dataset index (<root>/<dir>_game_counts.txt), random (game, move) sampling and
-- the queue_on_minibatch / do_queued_tasks pair (data.lua:29-37,53-96).  Without the `threads` rock the jobs run
-- inline when the queue is drained; callbacks still run on the main Lua state, as in the reference.
require 'dataloader'

Dataset = {}
Dataset.root = "/home/ubuntu/ebs_disk"
Dataset.directory = "train"

local queued = {}

function Dataset:new(dict)
    local result = {}
    for k, v in pairs(self) do result[k] = v end
    for k, v in pairs(dict or {}) do result[k] = v end
    self.__index = self
    return setmetatable(result, self)
end

function Dataset:init()
    self.games = {}
    local file = self.root .. "/" .. self.directory .. "_game_counts.txt"
    local counts = assert(io.open(file), "unable to find counts (probably bad dataset path)")
    for line in counts:lines() do
        local name, size = line:match("^(.-)\t%s*(%d+)%s*$")
        size = tonumber(size)
        if name and size > 0 then table.insert(self.games, {name = name, size = size}) end   -- empty games dropped
    end
    counts:close()
    self.initialized = true
end

function Dataset:generate_random_filename()
    if not self.initialized then self:init() end
    local game = self.games[math.random(1, #self.games)]
    return game.name .. "/" .. math.random(1, game.size)
end

function queue_on_minibatch(f, dataset, n)
    if dataset.initialized ~= true then dataset:init() end
    local filenames = {}
    for i = 1, n do filenames[i] = dataset:generate_random_filename() end
    table.insert(queued, {filenames = filenames, callback = f})
end

function do_queued_tasks()
    while #queued > 0 do
        local job = table.remove(queued, 1)
        job.callback(load_minibatch(job.filenames))
    end
end

-- Packed corpus (no counterpart in the reference, which keeps one Torch7 file per position): Corpus.pack(filenames, path)
-- once, then Corpus.open(path):raw_minibatch(n) returns the arguments of dg_model_set_batch_raw -- n memcpys out of one
-- mapped file instead of n torch.load calls.
local ffi = require 'ffi'
local deepgo = require 'deepgo'
Corpus = {}
Corpus.__index = Corpus

function Corpus.pack(filenames, path, threads)
    local arr = ffi.new("const char *[?]", #filenames)
    local keep = {}
    for i, f in ipairs(filenames) do keep[i] = ffi.new("char[?]", #f + 1, f); arr[i - 1] = keep[i] end
    deepgo.check(deepgo.lib.dg_corpus_pack(arr, #filenames, path, threads or 8))
    return path
end

function Corpus.open(path)
    local h = ffi.new("dg_corpus *[1]")
    deepgo.check(deepgo.lib.dg_corpus_open(path, h))
    local n = ffi.new("long long[1]")
    deepgo.check(deepgo.lib.dg_corpus_count(h[0], n))
    return setmetatable({h = ffi.gc(h[0], deepgo.lib.dg_corpus_close), count = tonumber(n[0])}, Corpus)
end

function Corpus:raw_minibatch(n)
    local idx = ffi.new("long long[?]", n)
    for i = 0, n - 1 do idx[i] = math.random(0, self.count - 1) end
    local b = {n = n, raw = ffi.new("uint8_t[?]", n * 9 * 361), player = ffi.new("uint8_t[?]", n), rank = ffi.new("uint8_t[?]", n),
               x = ffi.new("uint8_t[?]", n), y = ffi.new("uint8_t[?]", n)}
    deepgo.check(deepgo.lib.dg_corpus_read_batch(self.h, idx, n, b.raw, b.player, b.rank, b.x, b.y, 1))
    return b
end
