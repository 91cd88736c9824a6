"""B200-native pairwise descriptor matching (the one hot path of daovietanh190499/3D_Reconstruction).

The directory name starts with a digit, so import it with
    importlib.import_module("3d_reconstruction_b200")
The compute path is hand-written sm_100a CUDA behind the C-ABI in include/b200m.h; there is no CPU fallback.
"""
from ._lib import B200MError, LIB_PATH, SYMBOLS, get_option, match_stats, profile_read, set_option  # noqa: F401
from .bank import DescriptorBank, FloatBank, to_u8_descriptors  # noqa: F401
from .matcher import (B200Matcher, FlatMatches, MatchConfig, PairMatches, digest_matches, knn2_pairs, knn_match,  # noqa: F401
                      match, match_float, match_pairs, match_pairs_float, match_pairs_host)
from .assignment import AssignmentHead, assign_pairs  # noqa: F401
from .geometry import PoseBatch, essential_ransac  # noqa: F401
from .features import build_codebook, extract_features, sift_features  # noqa: F401
from .driver import Gates, bow_view_graph, match_view_graph, run_matching_stage, save_matches, vq_words  # noqa: F401
from .scheduler import (blocked_order, exhaustive_pairs, partition, partition_by_owner, window_pairs,  # noqa: F401
                        window_shard)
