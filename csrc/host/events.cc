// TensorBoard event-file writer: TFRecord framing + hand-encoded
// Event / Summary / SessionLog protos.  Replaces tf.summary.FileWriter + scalar/merge_all
// (reference worker.py:92-96,139) and the Supervisor's step/sec + session-log records.
//
// record = u64 length | u32 masked_crc32c(length) | payload | u32 masked_crc32c(payload)
#include "host_common.h"

#include <stdio.h>
#include <string.h>

#include <mutex>

using namespace dtfe_host;

namespace {

struct EventWriter {
  FILE* f = nullptr;
  std::mutex mu;
  uint64_t records = 0;
};

std::string EncodeDoubleField(int field, double v) {
  std::string s;
  put_tag(s, field, 1);
  uint64_t bits;
  memcpy(&bits, &v, 8);
  put_fixed64(s, bits);
  return s;
}

std::string EventPrefix(double wall_time, int64_t step) {
  std::string e = EncodeDoubleField(1, wall_time);
  if (step != 0) put_varint_field(e, 2, static_cast<uint64_t>(step));
  return e;
}

int WriteRecord(EventWriter* w, const std::string& payload) {
  std::string rec;
  uint64_t len = payload.size();
  char lenb[8];
  for (int i = 0; i < 8; ++i) lenb[i] = static_cast<char>((len >> (8 * i)) & 0xff);
  rec.append(lenb, 8);
  put_fixed32(rec, crc32c_mask(crc32c(lenb, 8)));
  rec.append(payload);
  put_fixed32(rec, crc32c_mask(crc32c(payload.data(), payload.size())));
  std::lock_guard<std::mutex> g(w->mu);
  if (!w->f) {
    set_error("event writer is closed");
    return -1;
  }
  if (fwrite(rec.data(), 1, rec.size(), w->f) != rec.size()) {
    set_error("short write to event file");
    return -1;
  }
  ++w->records;
  return 0;
}

}  // namespace

DTFE_HOST_API void* dtfe_events_open(const char* path) {
  FILE* f = fopen(path, "ab");
  if (!f) {
    set_error("cannot open event file %s", path);
    return nullptr;
  }
  EventWriter* w = new EventWriter();
  w->f = f;
  return w;
}

// First record of every file: Event{wall_time, file_version="brain.Event:2"}
DTFE_HOST_API int dtfe_events_write_version(void* h, double wall_time) {
  std::string e = EncodeDoubleField(1, wall_time);
  put_bytes_field(e, 3, "brain.Event:2");
  return WriteRecord(static_cast<EventWriter*>(h), e);
}

// Event{wall_time, step, summary{ value{tag, simple_value} ... }}
DTFE_HOST_API int dtfe_events_write_scalars(void* h, double wall_time, int64_t step, const char* const* tags,
                                            const float* values, int n) {
  std::string summary;
  for (int i = 0; i < n; ++i) {
    std::string val;
    put_bytes_field(val, 1, tags[i]);
    put_tag(val, 2, 5);
    uint32_t bits;
    memcpy(&bits, &values[i], 4);
    put_fixed32(val, bits);
    put_bytes_field(summary, 1, val);
  }
  std::string e = EventPrefix(wall_time, step);
  put_bytes_field(e, 5, summary);
  return WriteRecord(static_cast<EventWriter*>(h), e);
}

// n events at once (one per training step of a multi-step kernel launch): step[i] with the `ntags` scalars
// values[i*ntags .. ], all stamped `wall_time`.  Same bytes as n calls of dtfe_events_write_scalars.
DTFE_HOST_API int dtfe_events_write_scalars_batch(void* h, double wall_time, const int64_t* steps,
                                                  const char* const* tags, const float* values, int n, int ntags) {
  for (int i = 0; i < n; ++i)
    if (int rc = dtfe_events_write_scalars(h, wall_time, steps[i], tags, values + (size_t)i * ntags, ntags)) return rc;
  return 0;
}

// The per-step loss/accuracy summaries of one multi-step kernel launch, read straight from the arrays the GPU
// results were copied into (int32 observed global step, fp32 loss, fp32 accuracy per step): no per-value
// marshalling on the Python side.  Same bytes as dtfe_events_write_scalars(step[i], {tag0: loss[i], tag1: acc[i]}).
DTFE_HOST_API int dtfe_events_write_pairs_i32(void* h, double wall_time, const int32_t* steps, const float* v0,
                                              const float* v1, int n, const char* tag0, const char* tag1) {
  const char* tags[2] = {tag0, tag1};
  for (int i = 0; i < n; ++i) {
    const float vals[2] = {v0[i], v1[i]};
    if (int rc = dtfe_events_write_scalars(h, wall_time, steps[i], tags, vals, 2)) return rc;
  }
  return 0;
}

// Event{wall_time, step, session_log{status, checkpoint_path, msg}}; status: 1 START, 2 STOP, 3 CHECKPOINT
DTFE_HOST_API int dtfe_events_write_session_log(void* h, double wall_time, int64_t step, int status,
                                                const char* checkpoint_path, const char* msg) {
  std::string sl;
  if (status != 0) put_varint_field(sl, 1, static_cast<uint64_t>(status));
  if (checkpoint_path && *checkpoint_path) put_bytes_field(sl, 2, checkpoint_path);
  if (msg && *msg) put_bytes_field(sl, 3, msg);
  std::string e = EventPrefix(wall_time, step);
  put_bytes_field(e, 7, sl);
  return WriteRecord(static_cast<EventWriter*>(h), e);
}

// Event{wall_time, graph_def=bytes}
DTFE_HOST_API int dtfe_events_write_graph(void* h, double wall_time, const void* graph_def, size_t n) {
  std::string e = EncodeDoubleField(1, wall_time);
  put_bytes_field(e, 4, std::string(static_cast<const char*>(graph_def), n));
  return WriteRecord(static_cast<EventWriter*>(h), e);
}

DTFE_HOST_API int dtfe_events_flush(void* h) {
  EventWriter* w = static_cast<EventWriter*>(h);
  std::lock_guard<std::mutex> g(w->mu);
  if (w->f && fflush(w->f) != 0) {
    set_error("fflush failed");
    return -1;
  }
  return 0;
}

DTFE_HOST_API uint64_t dtfe_events_record_count(void* h) { return static_cast<EventWriter*>(h)->records; }

DTFE_HOST_API void dtfe_events_close(void* h) {
  EventWriter* w = static_cast<EventWriter*>(h);
  {
    std::lock_guard<std::mutex> g(w->mu);
    if (w->f) fclose(w->f);
    w->f = nullptr;
  }
  delete w;
}
