"""symmetry_b200 -- B200-native batched secretstream relay for Symmetry providers.

Only the one hot path of the reference is here: seal/open of streamed
chat-completion frames for all peer connections at once
(/root/reference/src/provider.ts:240-257), behind the C ABI in
include/symrelay.h.  The CUDA library is mandatory (no CPU fallback)."""
from .relay import (DELTA_DTYPE, DESC_DTYPE, F_EXTRACT, F_OPEN, PROVIDER_DEFAULT, PROVIDER_LLAMACPP,  # noqa: F401
                    PROVIDER_OLLAMA, ST_BAD_MAC, ST_OK, TILE_DTYPE, RelayError, SymRelay, provider_kind)

__all__ = ["SymRelay", "RelayError", "DESC_DTYPE", "TILE_DTYPE", "DELTA_DTYPE", "F_OPEN", "F_EXTRACT",
           "provider_kind"]
