// Internal constants shared by the kernels and the host-side module.
#ifndef PB2_INTERNAL_H
#define PB2_INTERNAL_H

// Return codes: same numeric values as the reference's PARSEC_SUCCESS / PARSEC_ERR_*
// (parsec/include/parsec/constants.h:13-25) so that a shim can forward them unchanged.
#define PB2_SUCCESS 0
#define PB2_ERROR (-1)
#define PB2_ERR_OUT_OF_RESOURCE (-2)
#define PB2_ERR_NOT_FOUND (-3)
#define PB2_ERR_BAD_PARAM (-4)
#define PB2_ERR_EXISTS (-5)
#define PB2_ERR_NOT_IMPLEMENTED (-6)
#define PB2_ERR_NOT_SUPPORTED (-7)
#define PB2_ERR_VALUE_OUT_OF_BOUNDS (-8)
#define PB2_ERR_DEVICE (-10)
#define PB2_ERR_CUDA PB2_ERR_DEVICE

// Hook return codes (parsec/runtime.h:140-146).
#define PB2_HOOK_RETURN_DONE 0
#define PB2_HOOK_RETURN_AGAIN (-1)
#define PB2_HOOK_RETURN_NEXT (-2)
#define PB2_HOOK_RETURN_DISABLE (-3)
#define PB2_HOOK_RETURN_ASYNC (-4)
#define PB2_HOOK_RETURN_ERROR (-5)

#endif
