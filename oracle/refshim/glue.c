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
void parsec_arena_release(parsec_data_copy_t* copy) { (void)copy; }

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
