-- BASELINE config 3: the paper's largest net, 12 layers x 192 filters, batch 256
require 'experiments'
experiment = basicGoExperiment:new{name = "12 layers x 192", useCuda = true, numLayers = 12, channelSize = 192, batchSize = 256,
                                   kernels = {5}, strides = {1}, channels = {37}, rate = .01, rateDecay = 1e-7}
experiment:run{iters = tonumber(arg and arg[1]) or 1000}
