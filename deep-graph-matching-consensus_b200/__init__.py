"""deep-graph-matching-consensus_b200 — B200-native DGMC hot path.

Put this directory on sys.path and `import dgmc`: the package mirrors the reference's public API
(dgmc.models.DGMC / RelCNN / SplineCNN / GIN / MLP, dgmc.utils.PairDataset / ValidPairDataset)
and runs the correspondence-and-consensus path on the kernels in csrc/ through the C ABI of
include/dgmc_b200.h (lib/libdgmc_b200.so, built by __graft_entry__.build()).
"""
