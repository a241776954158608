-- optimizer.lua of deep-go, re-hosted: SGD.new(rate, decay), SGD:step(parameters, grad).
-- The axpy runs on the device (dg_sgd_step: p += -rate*g, optimizer.lua:25); the multiplicative rate decay stays
-- here and is applied after use (optimizer.lua:26).  No momentum, no weight decay -- as in the reference.
local dg = require 'deepgo'

SGD = {}
SGD.__index = SGD

function SGD.new(rate, decay)
    return setmetatable({rate = rate, rate_decay = decay}, SGD)
end

function SGD:step(parameters, grad)
    assert(parameters.model == grad.model, "parameters and grad belong to different models")
    dg.check(dg.lib.dg_sgd_step(parameters.model.handle, self.rate))
    self.rate = self.rate * (1 - self.rate_decay)
end
