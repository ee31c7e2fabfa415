/* xvkaldi.h -- C ABI of the Kaldi ark/scp streaming I/O that sits either side of the embedding path
 * (SURVEY.md section 8(f) N1).  Host-only (no CUDA): lib/libxvkaldi.so.
 *
 * Replaces, function for function, the pure-Python reader/writer the reference extractor calls
 * (dataset/kaldi_io.py; callers egs/voxceleb/v1/nnet/lib/extract.py:63-64,93):
 *
 *   open_or_fd            dataset/kaldi_io.py:625-654    -> rspecifier / wspecifier parsing in xvk_reader_open / xvk_writer_open
 *   popen (pipes)         dataset/kaldi_io.py:657-690    -> "cmd |" and "| cmd" specifiers (exit status > 0 is an error)
 *   read_key              dataset/kaldi_io.py:693-707    -> xvk_next
 *   read_vec_flt[_ark]    dataset/kaldi_io.py:838-886    -> xvk_next + xvk_read_f32 / xvk_read_f64 (kind XVK_VEC_*)
 *   write_vec_flt         dataset/kaldi_io.py:915-946    -> xvk_write_vec_f32 / _f64
 *   read_mat_scp          dataset/kaldi_io.py:954-972    -> xvk_scp_open / xvk_scp_read_f32 / xvk_scp_read_many_f32
 *   read_mat_ark/read_mat dataset/kaldi_io.py:974-1011   -> xvk_next + xvk_read_f32
 *   _read_mat_binary      dataset/kaldi_io.py:1014-1031  -> FM / DM bodies
 *   _read_submat_binary   dataset/kaldi_io.py:1034-1053  -> xvk_read_rows_f32
 *   _read_mat_ascii       dataset/kaldi_io.py:1056-1068  -> text bodies (" [ ... ]")
 *   _read_compressed_mat  dataset/kaldi_io.py:1071-1115  -> CM bodies (per-column uint16 percentiles + uint8 piecewise-linear)
 *   write_mat             dataset/kaldi_io.py:1175-1208  -> xvk_write_mat_f32 / _f64
 *
 * Wire format (all little endian):  key ' ' '\0' 'B' token ...
 *   "FM " | "DM "   '\4' int32 rows '\4' int32 cols   rows*cols float32 | float64, row major
 *   "FV " | "DV "   '\4' int32 dim                    dim float32 | float64
 *   "CM "           float32 min, float32 range, int32 rows, int32 cols, cols x 4 uint16 percentiles, cols*rows uint8 (column major)
 * "CM2"/"CM3" are refused like the reference does (dataset/kaldi_io.py:1076).
 *
 * Every function returns XVK_OK (0) / a positive "have data" flag, or a negative XVK_E* code; the text of the last
 * error of a handle is xvk_reader_error / xvk_scp_error / xvk_writer_error (open errors: the err buffer).
 */
#ifndef XVKALDI_H_
#define XVKALDI_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define XVK_OK 0
#define XVK_EOPEN (-1)     /* cannot open file / start command */
#define XVK_EFORMAT (-2)   /* bad or unsupported wire format (message names the token, like UnknownMatrixHeader) */
#define XVK_EIO (-3)       /* short read / write error / command returned a non-zero status */
#define XVK_EINVAL (-4)
#define XVK_ECAPACITY (-5) /* destination too small */
#define XVK_ESTATE (-6)    /* call order (e.g. xvk_read_f32 without a pending entry) */

/* kinds reported by xvk_next */
#define XVK_MAT_F32 1
#define XVK_MAT_F64 2
#define XVK_MAT_CM 3
#define XVK_MAT_TEXT 4
#define XVK_VEC_F32 5
#define XVK_VEC_F64 6
#define XVK_VEC_TEXT 7

typedef struct xvk_reader xvk_reader;
typedef struct xvk_scp xvk_scp;
typedef struct xvk_writer xvk_writer;

/* ---- sequential ark reader ------------------------------------------------------------------------------------
 * rspecifier: [ark[,opts]:]path | path:offset | path.gz | "command |"   (open_or_fd).  With a byte offset the stream
 * starts at the "\0B" of one object (an scp target): open with keyless = 1 and the first xvk_next returns key "". */
xvk_reader* xvk_reader_open(const char* rspecifier, int keyless, char* err, size_t errcap);
/* Next object header.  1 = *key (valid until the next call), *kind, *rows (1 for vectors), *cols are set and the
 * body is pending; 0 = end of stream; < 0 error.  An unread pending body is skipped. */
int xvk_next(xvk_reader* r, const char** key, int32_t* kind, int32_t* rows, int32_t* cols);
/* Decode the pending body as fp32, row stride `ld` elements (ld >= cols).  FM: straight read into dst; DM: cast;
 * CM: table lookup, evaluated in float32 in the reference's operation order; text: strtod + cast. */
int xvk_read_f32(xvk_reader* r, float* dst, int64_t ld);
/* Same, as fp64 (what the reference returns for DM / DV objects). */
int xvk_read_f64(xvk_reader* r, double* dst, int64_t ld);
/* Rows [start, start+length) of the pending matrix (the reference's FeatureReader chunk read). */
int xvk_read_rows_f32(xvk_reader* r, int32_t start, int32_t length, float* dst, int64_t ld);
/* Packed batch for xv_forward: reads whole matrices back to back into dst [cap_rows, dim] while they fit
 * (rows, max_utts, key bytes); lengths[i] = rows of utterance i; keys = NUL-terminated strings back to back.
 * 1 = more objects may follow, 0 = end of stream reached; an object that does not fit stays pending. */
int xvk_read_batch_f32(xvk_reader* r, float* dst, int64_t cap_rows, int32_t dim, int32_t max_utts, int32_t* lengths,
                       char* keys, size_t keys_cap, int32_t* n_utts, int64_t* n_rows);
/* Regular (seekable, uncompressed) files only: xvk_read_batch_f32 scans the headers of a batch, then decodes the
 * bodies with `threads` workers through per-thread descriptors.  Streams and gzip stay sequential. */
int xvk_reader_set_threads(xvk_reader* r, int32_t threads);
const char* xvk_reader_error(const xvk_reader* r);
/* Closes the stream; for a command, waits for it: XVK_EIO if it returned a status > 0 (SubprocessFailed). */
int xvk_reader_close(xvk_reader* r);

/* ---- scp (random access) ----------------------------------------------------------------------------------------
 * Lines "key rxfile" with rxfile = path[:offset] (dataset/kaldi_io.py:966-969). */
xvk_scp* xvk_scp_open(const char* path, char* err, size_t errcap);
int64_t xvk_scp_count(const xvk_scp* s);
const char* xvk_scp_key(const xvk_scp* s, int64_t i);
const char* xvk_scp_rxfile(const xvk_scp* s, int64_t i);
/* Header of entry i (opens / seeks, reads the header only). */
int xvk_scp_dims(xvk_scp* s, int64_t i, int32_t* kind, int32_t* rows, int32_t* cols);
int xvk_scp_read_f32(xvk_scp* s, int64_t i, float* dst, int64_t ld);
/* Entries idx[0..n) decoded by `threads` workers (each with its own descriptors) into dst at row offsets
 * row_off[k] (row stride ld); rows_out[k] receives the row count.  cap_rows bounds dst. */
int xvk_scp_read_many_f32(xvk_scp* s, const int64_t* idx, int64_t n, float* dst, int64_t ld, const int64_t* row_off,
                          int64_t cap_rows, int32_t* rows_out, int32_t threads);
/* Rows [start[k], start[k] + length[k]) of entries idx[k], k in [0, n), decoded by `threads` workers into dst at row offsets
 * row_off[k]: the random segments of a training batch (KaldiDataRandomQueue's FeatureReader.read_segment calls,
 * dataset/data_loader.py:336-345 over dataset/kaldi_io.py:1034-1053) in one call -- only the bytes of the segments are read
 * (FM / DM; a compressed matrix is decoded row range by row range).  A segment that does not lie inside its matrix is
 * XVK_EINVAL ("The number of frames is not enough for length ..."). */
int xvk_scp_read_segments_f32(xvk_scp* s, const int64_t* idx, const int32_t* start, const int32_t* length, int64_t n, float* dst,
                              int64_t ld, const int64_t* row_off, int64_t cap_rows, int32_t threads);
/* Row / column counts of entries idx[0..n) (headers only), `threads` workers. */
int xvk_scp_dims_many(xvk_scp* s, const int64_t* idx, int64_t n, int32_t* rows_out, int32_t* cols_out, int32_t threads);
const char* xvk_scp_error(const xvk_scp* s);
void xvk_scp_close(xvk_scp* s);

/* ---- writer -----------------------------------------------------------------------------------------------------
 * wspecifier: [ark[,opts]:]path | path.gz | "| command" | ark,scp:ark_path,scp_path (the scp gets "key ark_path:offset"
 * lines, offset = position of the "\0B").  key "" writes a bare object (write_mat(filename, mat)). */
xvk_writer* xvk_writer_open(const char* wspecifier, char* err, size_t errcap);
int xvk_write_mat_f32(xvk_writer* w, const char* key, const float* m, int32_t rows, int32_t cols, int64_t ld);
int xvk_write_mat_f64(xvk_writer* w, const char* key, const double* m, int32_t rows, int32_t cols, int64_t ld);
int xvk_write_vec_f32(xvk_writer* w, const char* key, const float* v, int32_t dim);
int xvk_write_vec_f64(xvk_writer* w, const char* key, const double* v, int32_t dim);
/* n vectors of one matrix [n, ld] (the embeddings of a batch); keys = NUL-terminated strings back to back;
 * rows whose skip[i] != 0 are not written (utterances below min_chunk_size). */
int xvk_write_vecs_f32(xvk_writer* w, const char* keys, int64_t n, const float* data, int32_t dim, int64_t ld,
                       const uint8_t* skip);
int xvk_writer_flush(xvk_writer* w);
const char* xvk_writer_error(const xvk_writer* w);
int xvk_writer_close(xvk_writer* w);

/* CRC-32C (Castagnoli) of n bytes continuing from `crc` (0 to start): the checksum TensorFlow's tensor-bundle files
 * carry (tf_checkpoint.py). */
uint32_t xvk_crc32c(const void* data, size_t n, uint32_t crc);

const char* xvk_version(void);

#ifdef __cplusplus
}
#endif
#endif /* XVKALDI_H_ */
