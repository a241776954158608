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
-- eval() + optimizer:step (train.lua:111,131) as ONE device call: the captured step graph, with the gradient exchange inside it when
-- the model spans several GPUs; returns the cost eval() would have returned
function Model:train_step(mb, rate)
    stage(self, mb)
    local loss = ffi.new('float[1]')
    dg.check(lib.dg_train_step(self.handle, rate, loss))
    return loss[0]
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
    self.opt = {numLayers = numLayers, channelSize = channels[2], kFirst = kernels[1], precision = opt.precision, maxBatch = opt.maxBatch or 2048}
    self.ctx = dg.context(opt.device or default_device or 0)
    self.handle = dg.model(self.ctx, numLayers, channels[2], kernels[1], opt.precision, opt.maxBatch or 2048)
    local n = ffi.new('size_t[1]')
    dg.check(lib.dg_model_param_count(self.handle, n))
    self.P = tonumber(n[0])
    return self
end

-- experiments.lua:155-168: for nGPU > 1 the model becomes nGPU replicas, one per device of this process, behind the same handle
-- (nn.DataParallelTable(1): batches split on the batch dimension, gradients summed over NVLink peer memory inside the library);
-- the parameters are carried over.  nGPU == 1 returns the model unchanged, as the reference does.
function makeDataParallel(model, nGPU)
    if nGPU > 1 then
        local count = ffi.new('int[1]')
        dg.check(lib.dg_device_count(count))
        assert(nGPU <= count[0], 'number of GPUs less than nGPU specified')
        local buf = ffi.new('float[?]', model.P)
        dg.check(lib.dg_model_get_params(model.handle, buf, model.P))
        local o = model.opt
        local multi = setmetatable({}, getmetatable(model))
        multi.opt = o
        multi.ctx = dg.context_multi(nGPU)
        multi.handle = dg.model(multi.ctx, o.numLayers, o.channelSize, o.kFirst, o.precision, o.maxBatch)
        multi.P = model.P
        dg.check(lib.dg_model_set_params(multi.handle, buf, multi.P))
        model = multi
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

-- experiment:save / Experiment:load (experiments.lua:124-131).  One checkpoint format for every host of the library
-- (dg_checkpoint_save / dg_checkpoint_read: "DGCKPT02", key=value lines, "end", P float32 in Torch7 getParameters order); a file the
-- reference wrote with torch.save is read with dg_t7_read_model.
function Experiment:save(opt_filename)
    local filename = opt_filename or (self.id .. '.model')
    local costs = {}
    for i, v in ipairs(self.validation_costs) do costs[i] = string.format("%.17g", v) end
    local meta = string.format("name=%s\nrate=%.17g\nrateDecay=%.17g\nbatchSize=%d\nvalidationSize=%d\nvalidation_interval=%d\nnumGPUs=%d\n" ..
                               "id=%s\niterations=%d\noptimizer_rate=%.17g\noptimizer_rate_decay=%.17g\nvalidation_costs=%s\ndata_root=%s\n",
                               self.name or "", self.rate, self.rateDecay, self.batchSize, self.validationSize, self.validation_interval, self.numGPUs,
                               tostring(self.id), self.iterations, self.optimizer.rate, self.optimizer.rate_decay, table.concat(costs, ","), self.data_root)
    dg.check(lib.dg_checkpoint_save(self.model.handle, filename, meta))
end

function Experiment:load(filename)
    local f = assert(io.open(filename, 'rb'))
    local native = f:read(9) == "DGCKPT02\n"
    f:close()
    local n = ffi.new('size_t[1]')
    if not native then
        -- the reference's own checkpoint: flat vector, rate and iteration count; options from this prototype
        local rate, decay, iters = ffi.new('double[1]', -1), ffi.new('double[1]', -1), ffi.new('long long[1]', -1)
        dg.check(lib.dg_t7_read_model(filename, nil, 0, n, nil, nil, nil))
        local e = self:new()
        e:init()
        assert(tonumber(n[0]) == e.model.P, "the checkpoint's parameter count does not match the experiment's model")
        local buf = ffi.new('float[?]', e.model.P)
        dg.check(lib.dg_t7_read_model(filename, buf, e.model.P, n, rate, decay, iters))
        e.modelParameters:set(buf)
        if rate[0] >= 0 then e.optimizer = SGD.new(rate[0], decay[0] >= 0 and decay[0] or e.rateDecay) end
        if iters[0] >= 0 then e.iterations = tonumber(iters[0]) end
        return e
    end
    local meta = ffi.new('char[65536]')
    dg.check(lib.dg_checkpoint_read(filename, meta, 65536, nil, 0, n))
    local kv = {}
    for k, v in ffi.string(meta):gmatch("([%w_]+)=([^\n]*)\n") do kv[k] = v end
    local e = self:new{numLayers = tonumber(kv.numLayers), channelSize = tonumber(kv.channelSize), kernels = {tonumber(kv.kFirst)},
                       strides = {1}, channels = {tonumber(kv.inPlanes)}, rate = tonumber(kv.rate) or self.rate,
                       rateDecay = tonumber(kv.rateDecay) or self.rateDecay, batchSize = tonumber(kv.batchSize) or self.batchSize,
                       validationSize = tonumber(kv.validationSize) or self.validationSize,
                       validation_interval = tonumber(kv.validation_interval) or self.validation_interval,
                       numGPUs = tonumber(kv.numGPUs) or 1, name = kv.name, data_root = kv.data_root or self.data_root}
    e:init()
    local buf = ffi.new('float[?]', e.model.P)
    dg.check(lib.dg_checkpoint_read(filename, nil, 0, buf, e.model.P, n))
    e.modelParameters:set(buf)
    e.optimizer = SGD.new(tonumber(kv.optimizer_rate) or e.rate, tonumber(kv.optimizer_rate_decay) or e.rateDecay)
    e.iterations = tonumber(kv.iterations) or 0
    e.id = tonumber(kv.id) or kv.id
    e.validation_costs = {}
    for v in (kv.validation_costs or ""):gmatch("[^,]+") do table.insert(e.validation_costs, tonumber(v)) end
    return e
end
