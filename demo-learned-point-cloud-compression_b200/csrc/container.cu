// (6) Reference bitstream containers, host side: big-endian int32 / float64 fields around
// opaque byte strings.  No GPU work; lives in the same library so a C/C++ host needs nothing else to read
// or write what compress() / decompress() exchange.
#include <cstring>

#include "common.cuh"

namespace {

struct Writer {
  uint8_t* p;
  int64_t cap, pos;
  bool ok;
  void raw(const void* src, int64_t n) {
    if (n < 0 || pos + n > cap) {
      ok = false;
      return;
    }
    if (n) memcpy(p + pos, src, (size_t)n);
    pos += n;
  }
  void i32(int32_t v) {
    const uint32_t u = (uint32_t)v;
    const uint8_t b[4] = {(uint8_t)(u >> 24), (uint8_t)(u >> 16), (uint8_t)(u >> 8), (uint8_t)u};
    raw(b, 4);
  }
  void f64(double v) {
    uint64_t u;
    memcpy(&u, &v, 8);
    uint8_t b[8];
    for (int i = 0; i < 8; ++i) b[i] = (uint8_t)(u >> (56 - 8 * i));
    raw(b, 8);
  }
};

struct Reader {
  const uint8_t* p;
  int64_t n, pos;
  bool ok;
  bool need(int64_t k) {
    if (k < 0 || pos + k > n) ok = false;
    return ok;
  }
  int32_t i32() {
    if (!need(4)) return 0;
    const uint8_t* b = p + pos;
    pos += 4;
    return (int32_t)(((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | (uint32_t)b[3]);
  }
  double f64() {
    if (!need(8)) return 0.0;
    uint64_t u = 0;
    for (int i = 0; i < 8; ++i) u = (u << 8) | p[pos + i];
    pos += 8;
    double v;
    memcpy(&v, &u, 8);
    return v;
  }
  pcc_span_t span(int64_t k) {
    pcc_span_t s = {nullptr, 0};
    if (!need(k)) return s;
    s.data = p + pos;
    s.len = k;
    pos += k;
    return s;
  }
};

bool fits_i32(int64_t v) { return v >= 0 && v <= 0x7fffffffll; }

}  // namespace

extern "C" {

int64_t pcc_container_batched_bytes(int32_t num_frames, int64_t y_len, int64_t z_len, const pcc_span_t* points) {
  if (num_frames < 0 || y_len < 0 || z_len < 0 || (num_frames > 0 && !points)) return -1;
  int64_t total = 4 + 16 + 16 + y_len + z_len;
  for (int32_t i = 0; i < num_frames; ++i) {
    if (points[i].len < 0) return -1;
    total += 16 + points[i].len;
  }
  return total;
}

int pcc_container_write_batched(int32_t num_frames, double q_geo, double q_att, int32_t y_shape, int32_t z_shape,
                                pcc_span_t y, pcc_span_t z, const pcc_span_t* points, const int32_t* ks,
                                int32_t ks_ld, uint8_t* out, int64_t out_cap, int64_t* written) {
  PCC_CHECK_ARG(num_frames >= 0 && out && written && out_cap >= 0);
  PCC_CHECK_ARG(num_frames == 0 || (points && ks && ks_ld >= num_frames));
  PCC_CHECK_ARG(fits_i32(y.len) && fits_i32(z.len) && (y.len == 0 || y.data) && (z.len == 0 || z.data));
  Writer w = {out, out_cap, 0, true};
  w.i32(num_frames);
  w.f64(q_geo);
  w.f64(q_att);
  w.i32(y_shape);
  w.i32(z_shape);
  w.i32((int32_t)y.len);
  w.i32((int32_t)z.len);
  w.raw(y.data, y.len);
  w.raw(z.data, z.len);
  for (int32_t i = 0; i < num_frames; ++i) {
    PCC_CHECK_ARG(fits_i32(points[i].len) && (points[i].len == 0 || points[i].data));
    w.i32((int32_t)points[i].len);
    for (int l = 0; l < 3; ++l) w.i32(ks[(int64_t)l * ks_ld + i]);
    w.raw(points[i].data, points[i].len);
  }
  if (!w.ok) {
    pcc::set_error("pcc_container_write_batched: output buffer of %lld bytes is too small", (long long)out_cap);
    return PCC_E_WORKSPACE;
  }
  *written = w.pos;
  return PCC_OK;
}

int pcc_container_parse_batched(const uint8_t* data, int64_t nbytes, int32_t max_frames, int32_t* num_frames,
                                double* q_geo, double* q_att, int32_t* y_shape, int32_t* z_shape, pcc_span_t* y,
                                pcc_span_t* z, pcc_span_t* points, int32_t* ks, int32_t ks_ld) {
  PCC_CHECK_ARG(data && nbytes >= 0 && num_frames && q_geo && q_att && y_shape && z_shape && y && z);
  PCC_CHECK_ARG(max_frames >= 0 && (max_frames == 0 || (points && ks && ks_ld >= max_frames)));
  Reader r = {data, nbytes, 0, true};
  const int32_t nf = r.i32();
  *q_geo = r.f64();
  *q_att = r.f64();
  *y_shape = r.i32();
  *z_shape = r.i32();
  const int32_t y_len = r.i32(), z_len = r.i32();
  *y = r.span(y_len);
  *z = r.span(z_len);
  if (!r.ok || nf < 0 || nf > max_frames) {
    pcc::set_error("pcc_container_parse_batched: truncated blob or bad frame count (%d, room for %d)", nf, max_frames);
    return PCC_E_INVALID;
  }
  for (int32_t i = 0; i < nf; ++i) {
    const int32_t plen = r.i32();
    for (int l = 0; l < 3; ++l) ks[(int64_t)l * ks_ld + i] = r.i32();
    points[i] = r.span(plen);
    if (!r.ok) {
      pcc::set_error("pcc_container_parse_batched: truncated at frame %d", i);
      return PCC_E_INVALID;
    }
  }
  *num_frames = nf;
  return PCC_OK;
}

int64_t pcc_container_frames_bytes(int32_t num_frames, const pcc_span_t* y, const pcc_span_t* z,
                                   const pcc_span_t* points) {
  if (num_frames < 0 || (num_frames > 0 && (!y || !z || !points))) return -1;
  int64_t total = 4 + 16;
  for (int32_t i = 0; i < num_frames; ++i) {
    if (y[i].len < 0 || z[i].len < 0 || points[i].len < 0) return -1;
    total += 32 + y[i].len + z[i].len + points[i].len;
  }
  return total;
}

int pcc_container_write_frames(int32_t num_frames, double q_geo, double q_att, const int32_t* y_shapes,
                               const int32_t* z_shapes, const pcc_span_t* y, const pcc_span_t* z,
                               const pcc_span_t* points, const int32_t* ks, int32_t ks_ld, uint8_t* out,
                               int64_t out_cap, int64_t* written) {
  PCC_CHECK_ARG(num_frames >= 0 && out && written && out_cap >= 0);
  PCC_CHECK_ARG(num_frames == 0 || (y_shapes && z_shapes && y && z && points && ks && ks_ld >= num_frames));
  Writer w = {out, out_cap, 0, true};
  w.i32(num_frames);
  w.f64(q_geo);
  w.f64(q_att);
  for (int32_t i = 0; i < num_frames; ++i) {
    PCC_CHECK_ARG(fits_i32(points[i].len) && fits_i32(y[i].len) && fits_i32(z[i].len));
    PCC_CHECK_ARG((points[i].len == 0 || points[i].data) && (y[i].len == 0 || y[i].data) && (z[i].len == 0 || z[i].data));
    w.i32(y_shapes[i]);
    w.i32(z_shapes[i]);
    w.i32((int32_t)points[i].len);
    w.i32((int32_t)y[i].len);
    w.i32((int32_t)z[i].len);
    for (int l = 0; l < 3; ++l) w.i32(ks[(int64_t)l * ks_ld + i]);
    w.raw(points[i].data, points[i].len);
    w.raw(y[i].data, y[i].len);
    w.raw(z[i].data, z[i].len);
  }
  if (!w.ok) {
    pcc::set_error("pcc_container_write_frames: output buffer of %lld bytes is too small", (long long)out_cap);
    return PCC_E_WORKSPACE;
  }
  *written = w.pos;
  return PCC_OK;
}

int pcc_container_parse_frames(const uint8_t* data, int64_t nbytes, int32_t max_frames, int32_t* num_frames,
                               double* q_geo, double* q_att, int32_t* y_shapes, int32_t* z_shapes, pcc_span_t* y,
                               pcc_span_t* z, pcc_span_t* points, int32_t* ks, int32_t ks_ld) {
  PCC_CHECK_ARG(data && nbytes >= 0 && num_frames && q_geo && q_att);
  PCC_CHECK_ARG(max_frames >= 0 &&
                (max_frames == 0 || (y_shapes && z_shapes && y && z && points && ks && ks_ld >= max_frames)));
  Reader r = {data, nbytes, 0, true};
  const int32_t nf = r.i32();
  *q_geo = r.f64();
  *q_att = r.f64();
  if (!r.ok || nf < 0 || nf > max_frames) {
    pcc::set_error("pcc_container_parse_frames: truncated blob or bad frame count (%d, room for %d)", nf, max_frames);
    return PCC_E_INVALID;
  }
  for (int32_t i = 0; i < nf; ++i) {
    y_shapes[i] = r.i32();
    z_shapes[i] = r.i32();
    const int32_t plen = r.i32(), ylen = r.i32(), zlen = r.i32();
    for (int l = 0; l < 3; ++l) ks[(int64_t)l * ks_ld + i] = r.i32();
    points[i] = r.span(plen);
    y[i] = r.span(ylen);
    z[i] = r.span(zlen);
    if (!r.ok) {
      pcc::set_error("pcc_container_parse_frames: truncated at frame %d", i);
      return PCC_E_INVALID;
    }
  }
  *num_frames = nf;
  return PCC_OK;
}

}  // extern "C"
