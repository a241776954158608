-- experiments/repeated.lua of the reference on the LuaJIT host of this library: continue from the best previous checkpoint with
-- a fresh optimizer (repeated.lua:14-22).    luajit experiments/repeated.lua -gpu 1 -num 3 -iters 100
package.path = package.path .. ';../?.lua'
require 'experiments'
require 'optimizer'

local opt = {gpu = 1, num = 1, iters = 100, model = '0.95223742560484.model'}
for i = 1, #(arg or {}), 2 do
    local k = arg[i]:gsub('^%-', '')
    opt[k] = tonumber(arg[i + 1]) or arg[i + 1]
end

default_device = opt.gpu - 1                      -- cutorch.setDevice(opt.gpu) is 1-based
local experiment = Experiment:load(opt.model)
experiment.id = "repeatedRun" .. opt.num
experiment.optimizer = SGD.new(experiment.rate, experiment.rateDecay)

print("experiment rate", experiment.rate)

experiment:run{iters = opt.iters}
