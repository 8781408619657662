"""Which reference policies have a CUDA kernel in this build (names of
NewAlgorithmFactory, pkg/algorithm/types.go:26-47)."""
SUPPORTED_ALGOS = ("FIFO", "ElasticFIFO", "SRJF", "ElasticSRJF", "Tiresias", "ElasticTiresias",
                   "FfDLOptimizer", "AFS-L")
MUNKRES = True     # batched ComputeMunkresMax kernel (csrc/munkres.cu)
PLACEMENT = True   # bestFit + score + munkres kernel (csrc/place.cu)
