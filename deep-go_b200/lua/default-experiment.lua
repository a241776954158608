-- default-experiment.lua: 6 layers x 64 filters, batch 64, rate .512 (the reference's GPU configuration)
require 'experiments'

experiment = basicGoExperiment:new{
    name = "6 layers, rate=.512",
    group = "train",
    useCuda = true,
    validationSize = 1000,
    validation_interval = 20,
    numLayers = 6,
    channelSize = 64,
    batchSize = 64,
    kernels = {5}, strides = {1}, channels = {37},
    rate = .512,
    rateDecay = 1e-7,
    criterion = nn.ClassNLLCriterion()
}
experiment:init()
print('experiment id: ', experiment.id)
