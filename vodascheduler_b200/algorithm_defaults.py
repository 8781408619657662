"""Policy constants of the reference, as package variables (pkg/algorithm/tiresias.go:17-36,
elastic_tiresias.go:18-22) plus the two priority helpers (tiresias.go:109-119)."""
import math

TiresiasQueueNum = 2
TiresiasThresholdsSec = {0: 3600.0, 1: math.inf}
TiresiasPromoteKnob = 8
ElasticTiresiasCompactionThreshold = 10


def TiresiasDemotePriority(priority: int) -> int:
    return priority + 1 if priority < TiresiasQueueNum - 1 else priority


def TiresiasPromotePriority(priority: int) -> int:
    return 0
