/* stand-in for parsec/runtime.h (oracle/_ref build only) */
