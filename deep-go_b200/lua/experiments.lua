-- experiments.lua of deep-go, re-hosted: the Experiment prototype with the reference's option surface
-- (useCuda, numGPUs, directories, data_root, validationSize, validation_interval, batchSize, name, kernels,
-- strides, channels, numLayers, channelSize, rate, rateDecay, criterion), getBasicModel, makeDataParallel,
-- save/load.  The model is a handle on a dg_model that keeps the nn.Module call shape train.lua uses.
require 'optimizer'
require 'train'
require 'data'
local dg = require 'deepgo'
local ffi, lib = dg.ffi, dg.lib

nn = nn or {}
-- nn.ClassNLLCriterion() (experiments.lua:45): the loss and its gradient are fused into the device step; the
-- object only exposes the numbers in the reference's call order
function nn.ClassNLLCriterion()
    local c = {}
    function c:cuda() return self end
    function c:forward(preds, targets) return preds.model:loss() end
    function c:backward(preds, targets) return preds end
    return c
end

local Model = {}
Model.__index = Model

local function stage(self, mb)
    local labels = ffi.new('int32_t[?]', mb.size)
    for i = 0, mb.size - 1 do labels[i] = mb.output[i] - 1 end                 -- Lua side is 1-based (dataloader.lua:89)
    if mb.raw then   -- records are at hand: ship 3.2 KB/position and encode on the device
        dg.check(lib.dg_model_set_batch_raw(self.handle, mb.raw, mb.player, mb.rank, nil, mb.x, mb.y, mb.size))
    else
        dg.check(lib.dg_model_set_batch_planes(self.handle, mb.input, labels, mb.size))
    end
end

function Model:cuda() return self end
function Model:forward(mb)
    stage(self, mb)
    dg.check(lib.dg_forward(self.handle, nil))
    return {model = self, size = mb.size}
end
function Model:loss()
    local errs, cost = ffi.new('int[1]'), ffi.new('float[1]')
    dg.check(lib.dg_eval_batch(self.handle, errs, cost))
    self.last_errors = errs[0]
    return cost[0]
end
function Model:backward(mb, df_do)
    dg.check(lib.dg_loss_backward(self.handle, nil))
end
function Model:eval_slice(mb, targets, lo, hi)
    local n = hi - lo + 1
    local off = lo - 1
    local labels = ffi.new('int32_t[?]', n)
    for i = 0, n - 1 do labels[i] = targets[off + i] - 1 end
    dg.check(lib.dg_model_set_batch_planes(self.handle, mb.input + off * TOTAL * SIZE * SIZE, labels, n))
    local errs, cost = ffi.new('int[1]'), ffi.new('float[1]')
    dg.check(lib.dg_eval_batch(self.handle, errs, cost))
    return errs[0], cost[0]
end
function Model:getParameters()
    local model = self
    local params = {model = model}
    local grads = {model = model}
    function grads:zero() dg.check(lib.dg_zero_grads(model.handle)) end
    function params:get(buf) dg.check(lib.dg_model_get_params(model.handle, buf, model.P)) end
    function params:set(buf) dg.check(lib.dg_model_set_params(model.handle, buf, model.P)) end
    function grads:get(buf) dg.check(lib.dg_model_get_grads(model.handle, buf, model.P)) end
    return params, grads
end

-- numLayers x [pad, conv k, + per-channel bias, + per-point bias, ReLU] + log-softmax(361)  (experiments.lua:133-153)
function getBasicModel(numLayers, kernels, channels, opt)
    opt = opt or {}
    local self = setmetatable({}, Model)
    self.ctx = dg.context(opt.device or 0)
    self.handle = dg.model(self.ctx, numLayers, channels[2], kernels[1], opt.precision, opt.maxBatch or 2048)
    local n = ffi.new('size_t[1]')
    dg.check(lib.dg_model_param_count(self.handle, n))
    self.P = tonumber(n[0])
    return self
end

-- one process per GPU: the launcher starts numGPUs processes (torchrun-style env RANK / WORLD_SIZE) and every rank
-- binds the library's NCCL communicator; inside one process the model is returned unchanged (experiments.lua:155-168)
function makeDataParallel(model, nGPU)
    if nGPU > 1 then
        local count = ffi.new('int[1]')
        dg.check(lib.dg_device_count(count))
        assert(nGPU <= count[0], 'number of GPUs less than nGPU specified')
    end
    return model
end

Experiment = {
    useCuda = false, numGPUs = 1,
    directories = {train = "train", validation = "validation", test = "test"},
    datasets = {}, data_root = "/home/ubuntu/ebs_disk",
    validationSize = 2000, validation_interval = 2000, batchSize = 32
}

function Experiment:new(dict)
    local result = {}
    for k, v in pairs(self) do result[k] = v end
    for k, v in pairs(dict or {}) do result[k] = v end
    self.__index = self
    return setmetatable(result, self)
end

basicGoExperiment = Experiment:new{
    name = "basicGoExperiment", kernels = {5}, strides = {1}, channels = {37}, numLayers = 3, channelSize = 64,
    rate = .01, rateDecay = 1e-7, criterion = nn.ClassNLLCriterion()
}

function Experiment:init_data()
    self.datasets = {}
    for group, directory in pairs(self.directories) do
        self.datasets[group] = Dataset:new{root = self.data_root, directory = directory}
        self.datasets[group]:init()
    end
end

function Experiment:init()
    self.id = math.random()
    self.iterations = 0
    self.initialized = true
    self.optimizer = SGD.new(self.rate, self.rateDecay)
    self.validation_costs = {}
    self:init_data()
    local kernels, strides, channels = {self.kernels[1]}, {self.strides[1]}, {self.channels[1]}   -- per-instance copies
    for i = 2, self.numLayers do
        kernels[i], strides[i], channels[i] = 3, 1, self.channelSize
    end
    channels[#channels + 1] = 1
    self.kernels, self.strides, self.channels = kernels, strides, channels
    self.model = getBasicModel(self.numLayers, kernels, channels, {precision = self.precision, maxBatch = math.max(self.batchSize, 256)})
    if self.useCuda then
        self.model = makeDataParallel(self.model:cuda(), self.numGPUs)
    end
    self.modelParameters, self.grads = self.model:getParameters()
end

function Experiment:run(params)
    assert(params.iters > 0)
    if not self.initialized then self:init() end
    local start_time = os.clock()
    local train_cost, validation_cost = train(self, params)
    self.runningTime = os.clock() - start_time
    return train_cost, validation_cost
end

-- checkpoint = flat parameter vector (Torch7 order) + optimizer rate + iteration count (experiments.lua:124-131)
function Experiment:save(opt_filename)
    local filename = opt_filename or (self.id .. '.model')
    local buf = ffi.new('float[?]', self.model.P)
    self.modelParameters:get(buf)
    local f = assert(io.open(filename, 'wb'))
    f:write(string.format("deepgo-b200 %d %d %d %.17g %d\n", self.numLayers, self.channelSize, self.model.P, self.optimizer.rate, self.iterations))
    f:write(ffi.string(buf, self.model.P * 4))
    f:close()
end

function Experiment:load(filename)
    local f = assert(io.open(filename, 'rb'))
    local L, C, P, rate, iterations = f:read('*l'):match("deepgo%-b200 (%d+) (%d+) (%d+) (%S+) (%d+)")
    local e = self:new{numLayers = tonumber(L), channelSize = tonumber(C)}
    e:init()
    local bytes = f:read(tonumber(P) * 4)
    f:close()
    local buf = ffi.new('float[?]', tonumber(P))
    ffi.copy(buf, bytes, tonumber(P) * 4)
    e.modelParameters:set(buf)
    e.optimizer = SGD.new(tonumber(rate), e.rateDecay)
    e.iterations = tonumber(iterations)
    return e
end
