"""cmf_b200 -- B200-native (sm_100a) non-square log-det + g_ij hot path of Canonical Manifold
Flows, a drop-in behind the reference's ``--model non-square`` config surface."""
__version__ = "0.1.0"
