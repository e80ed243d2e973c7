/* pp_internal.h — private structures of the C99 host library (libpoporon.so).
 *
 * The host library owns: parameter validation, field/generator construction (cheap, once per
 * handle), the opaque config/handle/erasure objects of the public API, and the host<->device
 * staging pipeline.  Every codeword operation is executed by libpoporon_cuda.so. */
#ifndef PP_INTERNAL_H
#define PP_INTERNAL_H

#include <poporon.h>
#include <poporon_cuda.h>

#define PP_VERSION_ID  20000000u /* reference src/internal/common.h:29 */
#define PP_MAX_DEVICES 16
#define PP_SLOTS       4         /* staging slots (CUDA streams + buffers) per device: H2D of chunk i+2.. runs while chunk i drains */

#ifndef POPORON_BUILDTIME
#define POPORON_BUILDTIME 1
#endif

struct _poporon_gf_t {
    uint8_t symbol_size;
    uint16_t nn; /* 2^m - 1, also the log of zero */
    uint16_t poly;
    uint16_t alpha_to[256];
    uint16_t index_of[256];
};

typedef struct _poporon_rs_t {
    poporon_gf_t *gf;
    uint16_t fcr, prim, iprim;
    uint8_t nroots;
    uint16_t genpoly[256]; /* index form, nroots + 1 entries */
} poporon_rs_t;

typedef struct _poporon_bch_t {
    poporon_gf_t *gf;
    uint8_t t;
    uint16_t n, k, parity_bits, gen_deg;
    uint32_t gen;
    struct _poporon_t *self; /* lazily created handle backing the word-level API */
} poporon_bch_t;

struct _poporon_erasure_t {
    uint32_t capacity;
    uint32_t count;
    uint32_t *positions;
    uint16_t *corrections; /* never written; kept because the reference allocates it */
};

struct _poporon_config_t {
    poporon_fec_type_t fec_type;
    uint8_t symbol_size;
    uint16_t gfpoly;
    /* RS */
    uint16_t fcr, prim;
    uint8_t nroots;
    poporon_erasure_t *erasure;
    uint16_t *syndrome;
    /* BCH */
    uint8_t t;
};

/* one staging slot: a stream plus grow-only device buffers */
typedef struct {
    void *stream;
    void *kdone;       /* event: this slot's kernels have finished (the decode context's scratch is free again) */
    bool kdone_set;
    uint8_t *d_span;   /* data (and, for merged layouts, parity) bytes */
    size_t span_cap;
    uint8_t *d_pspan;  /* parity bytes when not merged */
    size_t pspan_cap;
    uint8_t *d_aux;    /* ok | corrected | positions | counts | syndromes */
    size_t aux_cap;
} pp_slot_t;

typedef struct {
    int device;
    bool ready;
    poporon_cuda_rs_t *rs;
    poporon_cuda_bch_t *bch;
    pp_slot_t slot[PP_SLOTS];
} pp_lane_t;

struct _poporon_t {
    poporon_fec_type_t fec_type;
    poporon_rs_t *rs;
    poporon_bch_t *bch;
    poporon_erasure_t *erasure; /* borrowed */
    uint16_t *ext_syndrome;     /* borrowed */
    size_t last_corrected;
    int32_t last_num_errors;
    size_t parity_bytes;        /* RS: nroots, BCH: ceil(parity_bits / 8) */
    size_t nlanes;
    pp_lane_t lane[PP_MAX_DEVICES];
    size_t chunk_codewords;     /* staging granularity */
};

/* rs_code.c / bch_code.c */
poporon_rs_t *poporon_rs_create(uint8_t symbol_size, uint16_t generator_polynomial, uint16_t first_consecutive_root,
                                uint16_t primitive_element, uint8_t num_roots);
void poporon_rs_destroy(poporon_rs_t *rs);
poporon_bch_t *poporon_bch_create(uint8_t symbol_size, uint16_t generator_polynomial, uint8_t t);
void poporon_bch_destroy(poporon_bch_t *bch);
uint16_t poporon_bch_get_codeword_length(const poporon_bch_t *bch);
uint16_t poporon_bch_get_data_length(const poporon_bch_t *bch);
uint8_t poporon_bch_get_correction_capability(const poporon_bch_t *bch);
bool poporon_bch_encode(poporon_bch_t *bch, uint32_t data, uint32_t *codeword);
bool poporon_bch_decode(poporon_bch_t *bch, uint32_t received, uint32_t *corrected, int32_t *num_errors);
uint32_t poporon_bch_extract_data(const poporon_bch_t *bch, uint32_t codeword);

/* handle.c */
poporon_t *pp_handle_from_bch(poporon_bch_t *bch_borrowed);
void pp_handle_release_lanes(poporon_t *h);

/* pipeline.c */
typedef struct {
    int op; /* 0 encode, 1 decode, 2 syndromes only (RS) */
    int mode;
    uint8_t *data;
    size_t data_stride, size;
    uint8_t *parity;
    size_t parity_stride;
    size_t n;
    const uint32_t *pos;
    size_t pos_stride;
    const uint32_t *cnt;
    const uint16_t *syn;
    uint8_t *ok;
    uint32_t *corrected;
    uint16_t *syn_out; /* op 2: [n][nroots] */
} pp_job_t;

bool pp_lane_prepare(poporon_t *h, pp_lane_t *lane);
bool pp_run_host_job(poporon_t *h, const pp_job_t *job);
bool pp_run_device_job(poporon_t *h, const pp_job_t *job, int device, void *stream, bool sync);

#endif /* PP_INTERNAL_H */
