/* Symbols the reference sources compiled by `make ref` need from the rest of the runtime, plus small accessors so
 * that a test can read and set fields of the reference's own parsec_data_t / parsec_data_copy_t without re-declaring
 * their layout.  Written for this repository; includes (does not copy) the reference headers. */
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "parsec/parsec_config.h"
#include "parsec/data_internal.h"
#include "parsec/mca/device/device.h"

/* --- the pieces of the runtime that data.c / grid_2Dcyclic.c reference ---------------------------------- */
uint32_t parsec_nb_devices = 4; /* host, "recursive", two GPUs */
int vpmap_get_nb_vp(void) { return 1; }
void parsec_debug_verbose(int level, int output_id, const char* fmt, ...) { (void)level; (void)output_id; (void)fmt; }
void parsec_warning(const char* fmt, ...) { (void)fmt; }
int parsec_mca_device_registration_completed(parsec_context_t* ctx) { (void)ctx; return 1; }
parsec_device_module_t* parsec_mca_device_get(uint32_t idx) { (void)idx; return NULL; }
void parsec_fatal(const char* fmt, ...) { fprintf(stderr, "parsec_fatal: %s\n", fmt); abort(); }

/* --- accessors ------------------------------------------------------------------------------------------ */
static char fake_dc[4096]; /* parsec_data_create only reads dc->default_dtt */

void* ref_data_new(uint64_t key) {
    static int inited = 0;
    if (!inited) {
        parsec_data_init(NULL);
        inited = 1;
    }
    parsec_data_t* holder = NULL;
    return parsec_data_create(&holder, (parsec_data_collection_t*)fake_dc, key, NULL, 1024, 0);
}
int ref_copy_attach_new(void* data, int device) { /* a fresh device copy: INVALID, version 0 (reserve, :1143-1151) */
    parsec_data_copy_t* c = parsec_data_copy_new((parsec_data_t*)data, (uint8_t)device, 0, 0);
    if (!c) return -1;
    c->coherency_state = PARSEC_DATA_COHERENCY_INVALID;
    c->version = 0;
    return 0;
}
int ref_copy_detach(void* data, int device) {
    parsec_data_t* d = (parsec_data_t*)data;
    parsec_data_copy_t* c = d->device_copies[device];
    return c ? parsec_data_copy_detach(d, c, (uint8_t)device) : -1;
}
int ref_start_transfer(void* data, int device, int access) {
    return parsec_data_start_transfer_ownership_to_copy((parsec_data_t*)data, (uint8_t)device, (uint8_t)access);
}
void ref_end_transfer(void* data, int device, int access) {
    parsec_data_end_transfer_ownership_to_copy((parsec_data_t*)data, (uint8_t)device, (uint8_t)access);
}
/* out: present, coherency, readers, version */
void ref_copy_state(void* data, int device, int* out) {
    parsec_data_copy_t* c = ((parsec_data_t*)data)->device_copies[device];
    out[0] = c != NULL;
    out[1] = c ? c->coherency_state : -1;
    out[2] = c ? c->readers : 0;
    out[3] = c ? (int)c->version : 0;
}
int ref_data_owner(void* data) { return ((parsec_data_t*)data)->owner_device; }
void ref_copy_set(void* data, int device, int coherency, int version, int readers) {
    parsec_data_copy_t* c = ((parsec_data_t*)data)->device_copies[device];
    if (!c) return;
    if (coherency >= 0) c->coherency_state = (parsec_data_coherency_t)coherency;
    if (version >= 0) c->version = (uint32_t)version;
    if (readers >= 0) c->readers = readers;
}

/* --- more of the runtime, needed once arena.c / parsec_hash_table.c / the matrix collection are compiled in ---- */
#include <stdbool.h>
static void* glue_alloc(size_t size) { return malloc(size); }
static void glue_free(void* p) { free(p); }
parsec_data_allocate_t parsec_data_allocate = glue_alloc; /* function-pointer globals of the runtime (runtime.h:100-101) */
parsec_data_free_t parsec_data_free = glue_free;
int parsec_mca_param_reg_int_name(const char* type, const char* name, const char* help, bool internal, bool read_only,
                                  int default_value, int* current_value) {
    (void)type; (void)name; (void)help; (void)internal; (void)read_only;
    if (current_value) *current_value = default_value;
    return 0;
}
int parsec_mca_param_lookup_int(int index, int* value) { (void)index; (void)value; return -1; }

/* --- the reference's 2D block-cyclic collection, driven through its own vtable ------------------------------ */
#include "parsec/data_dist/matrix/two_dim_rectangle_cyclic.h"

void* ref_bc_new(int myrank, int mb, int nb, int lm, int ln, int i, int j, int m, int n, int P, int Q, int kp, int kq,
                 int ip, int jq) {
    parsec_matrix_block_cyclic_t* dc = (parsec_matrix_block_cyclic_t*)calloc(1, sizeof(*dc));
    parsec_matrix_block_cyclic_init(dc, PARSEC_MATRIX_DOUBLE, PARSEC_MATRIX_TILE, myrank, mb, nb, lm, ln, i, j, m, n, P, Q,
                                    kp, kq, ip, jq);
    return dc;
}
void* ref_bc_lapack_new(int myrank, int mb, int nb, int lm, int ln, int i, int j, int m, int n, int P, int Q, int kp,
                        int kq, int ip, int jq, int mloc, int nloc) {
    parsec_matrix_block_cyclic_t* dc = (parsec_matrix_block_cyclic_t*)calloc(1, sizeof(*dc));
    parsec_matrix_block_cyclic_lapack_init(dc, PARSEC_MATRIX_DOUBLE, PARSEC_MATRIX_LAPACK, myrank, mb, nb, lm, ln, i, j, m, n,
                                           P, Q, kp, kq, ip, jq, mloc, nloc);
    return dc;
}
/* out: mt, nt, lmt, lnt, lm, ln, llm, lln, bsiz, nb_elem_r, nb_elem_c, nb_local_tiles, rrank, crank */
void ref_bc_geometry(void* p, int* out) {
    parsec_matrix_block_cyclic_t* dc = (parsec_matrix_block_cyclic_t*)p;
    int v[14] = {dc->super.mt, dc->super.nt, dc->super.lmt, dc->super.lnt, dc->super.lm, dc->super.ln, dc->super.llm,
                 dc->super.lln, dc->super.bsiz, dc->nb_elem_r, dc->nb_elem_c, dc->super.nb_local_tiles, dc->grid.rrank,
                 dc->grid.crank};
    memcpy(out, v, sizeof(v));
}
int ref_bc_rank_of(void* p, int m, int n) {
    parsec_data_collection_t* dc = (parsec_data_collection_t*)p;
    return (int)dc->rank_of(dc, m, n);
}
uint64_t ref_bc_data_key(void* p, int m, int n) {
    parsec_data_collection_t* dc = (parsec_data_collection_t*)p;
    return dc->data_key(dc, m, n);
}
/* byte offset of tile (m, n) inside the local matrix, its parsec_data_t key and nb_elts, through the real data_of */
int64_t ref_bc_tile_offset(void* p, int m, int n, uint64_t* key, uint32_t* nb_elts) {
    parsec_matrix_block_cyclic_t* bc = (parsec_matrix_block_cyclic_t*)p;
    parsec_data_collection_t* dc = (parsec_data_collection_t*)p;
    static char base[1]; /* symbolic: data_of only does pointer arithmetic on dc->mat */
    if (!bc->mat) bc->mat = base;
    parsec_data_t* d = dc->data_of(dc, m, n);
    if (key) *key = d->key;
    if (nb_elts) *nb_elts = d->nb_elts;
    return (int64_t)((char*)d->device_copies[0]->device_private - (char*)bc->mat);
}
