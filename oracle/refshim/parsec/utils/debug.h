/* stand-in for parsec/utils/debug.h (oracle/_ref build only): logging compiled out */
#ifndef PB2_REFSHIM_DEBUG_H
#define PB2_REFSHIM_DEBUG_H
void parsec_debug_verbose(int level, int output_id, const char* fmt, ...);
#define PARSEC_DEBUG_VERBOSE(...) do { } while (0)
#define parsec_debug_output 0
#endif
