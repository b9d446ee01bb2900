/* stand-in for parsec/vpmap.h (oracle/_ref build only) */
int vpmap_get_nb_vp(void);
