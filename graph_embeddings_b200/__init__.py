"""B200-native hot path of Phaken/graph-embeddings: co-occurrence (BCA) builder + GloVe/AdaGrad
trainer + embedding store, behind the reference's CoOccurrenceMatrix / IOptimizer seam.
All compute is in csrc/ (CUDA, sm_100a) behind the C ABI of include/graphemb.h."""
from .api import (Adagrad, Adam, AMSGrad, BCA, BookmarkColoring, Configuration, CoOccurrenceMatrix,  # noqa: F401
                  GraphEmbError, ImportedMatrix, InvalidConfigurationException, IOptimizer, Opt,
                  Optimum, createOptimizer)
from .config import load_config, parse_config  # noqa: F401,E402
from .writer import EmbeddingTextWriter, createFileName  # noqa: F401,E402
