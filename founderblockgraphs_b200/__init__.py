"""founderblockgraphs_b200 -- B200 (sm_100a) implementation of the founderblockgraphs
hot path (semi-repeat-free segmentation + founder-block construction) behind the
reference's CLI and a C ABI (include/fbg_cuda.h).  See DESIGN.md."""
from . import api, synth  # noqa: F401
from .api import (Context, FbgError, NoSegmentation, load_library, main, output_efg, read_input, segment,  # noqa: F401
                  segment2elasticValid, segment_elastic_minmaxlength)
