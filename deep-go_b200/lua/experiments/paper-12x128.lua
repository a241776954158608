-- BASELINE config 2: the paper-size policy net, 12 layers x 128 filters, batch 128, one B200
require 'experiments'
experiment = basicGoExperiment:new{name = "12 layers x 128", useCuda = true, numLayers = 12, channelSize = 128, batchSize = 128,
                                   kernels = {5}, strides = {1}, channels = {37}, rate = .01, rateDecay = 1e-7}
experiment:run{iters = tonumber(arg and arg[1]) or 1000}
