"""Host-side helpers of the B200-native rasterizer hot path (loader for the C-ABI
library, synthetic scenes, view-sharded data-parallel step)."""
