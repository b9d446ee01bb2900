/* Symbols the two reference sources need from the rest of the runtime (oracle/_ref build only). */
#include <stdarg.h>
int vpmap_get_nb_vp(void) { return 1; }
void parsec_debug_verbose(int level, int output_id, const char* fmt, ...) { (void)level; (void)output_id; (void)fmt; }
