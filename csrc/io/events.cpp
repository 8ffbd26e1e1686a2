// b200ps N8 — native TensorBoard event-file writer (the EventsWriter behind tf.train.SummaryWriter).
//
// example.py writes the scalars `cost` and `accuracy` EVERY step into ./logs/mnist/{job}_{task}
// (example.py:131-135,154,171), synchronously inside its hot loop.  Here add_scalars() only encodes the
// Event proto, frames it as a TFRecord and appends it to an in-memory queue; a writer thread owns the
// file, drains the queue and flushes at most every `flush_secs`, so a 13 us training step never waits
// for the file system.
//
// Record = u64 length | masked crc32c(length) | payload | masked crc32c(payload); the first record of a
// file is Event{wall_time, file_version: "brain.Event:2"}; a scalar record is
// Event{wall_time, step, summary{value{tag, simple_value}...}}.  b200ps/io/events.py holds the Python twin
// and the reader used by the tests.
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>

#include "../common.h"

using namespace b200ps;

extern "C" uint32_t b200ps_crc32c_extend(uint32_t init_crc, const void* data, uint64_t n);

namespace {

inline uint32_t ev_mask(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xa282ead8u; }

void ev_varint(std::string& s, uint64_t v) {
  while (v >= 0x80) { s.push_back((char)((v & 0x7F) | 0x80)); v >>= 7; }
  s.push_back((char)v);
}
void ev_tag(std::string& s, int field, int wt) { ev_varint(s, ((uint64_t)field << 3) | (uint64_t)wt); }
void ev_bytes(std::string& s, int field, const std::string& b) {
  ev_tag(s, field, 2);
  ev_varint(s, b.size());
  s += b;
}
void ev_raw(std::string& s, const void* p, size_t n) { s.append(static_cast<const char*>(p), n); }

std::string record(const std::string& payload) {
  std::string out;
  const uint64_t n = payload.size();
  ev_raw(out, &n, 8);
  const uint32_t c1 = ev_mask(b200ps_crc32c_extend(0, &n, 8));
  ev_raw(out, &c1, 4);
  out += payload;
  const uint32_t c2 = ev_mask(b200ps_crc32c_extend(0, payload.data(), payload.size()));
  ev_raw(out, &c2, 4);
  return out;
}

std::string event(double wall_time, int64_t step, const char* file_version, const std::string* summary) {
  std::string e;
  ev_tag(e, 1, 1);
  ev_raw(e, &wall_time, 8);                       // wall_time is always present
  if (step != 0) { ev_tag(e, 2, 0); ev_varint(e, (uint64_t)step); }
  if (file_version) ev_bytes(e, 3, file_version);
  if (summary) ev_bytes(e, 5, *summary);
  return e;
}

struct EventsWriter {
  FILE* f = nullptr;
  std::mutex mu;
  std::condition_variable cv;
  std::string pending;          // framed records not yet handed to the file
  bool stop = false, flush_now = false, writing = false;
  double flush_secs = 2.0;
  std::thread worker;
  uint64_t records = 0;

  void run() {
    std::unique_lock<std::mutex> lk(mu);
    for (;;) {
      cv.wait_for(lk, std::chrono::duration<double>(flush_secs), [&] { return stop || flush_now; });
      std::string take;
      take.swap(pending);
      const bool last = stop;
      flush_now = false;
      writing = true;
      lk.unlock();
      if (!take.empty()) fwrite(take.data(), 1, take.size(), f);
      fflush(f);
      lk.lock();
      writing = false;
      cv.notify_all();                            // wakes flush() callers
      if (last && pending.empty()) return;
    }
  }
};

double now_seconds() {
  return std::chrono::duration<double>(std::chrono::system_clock::now().time_since_epoch()).count();
}

}  // namespace

extern "C" int b200ps_events_open(void** out, const char* path, double flush_secs) {
  B200PS_REQUIRE(out && path && *path, "events_open: null argument");
  EventsWriter* w = new EventsWriter();
  w->f = fopen(path, "wb");
  if (!w->f) {
    delete w;
    return set_error(B200PS_ERR_INVALID, "events_open: cannot open %s", path);
  }
  w->flush_secs = flush_secs > 0 ? flush_secs : 2.0;
  const std::string first = record(event(now_seconds(), 0, "brain.Event:2", nullptr));
  fwrite(first.data(), 1, first.size(), w->f);
  fflush(w->f);
  w->worker = std::thread([w] { w->run(); });
  *out = w;
  return 0;
}

// One Event with n scalar values (tags[i], values[i]) at `step`; wall_time < 0 = now.  Never blocks on I/O.
extern "C" int b200ps_events_add_scalars(void* writer, double wall_time, int64_t step, const char* const* tags,
                                         const float* values, int32_t n) {
  EventsWriter* w = static_cast<EventsWriter*>(writer);
  B200PS_REQUIRE(w && n >= 0 && (n == 0 || (tags && values)), "events_add_scalars: bad arguments");
  std::string summary;
  for (int i = 0; i < n; ++i) {
    std::string value;
    if (tags[i] && *tags[i]) ev_bytes(value, 1, tags[i]);
    ev_tag(value, 2, 5);
    ev_raw(value, &values[i], 4);                 // simple_value, written even when 0
    ev_bytes(summary, 1, value);
  }
  const std::string rec = record(event(wall_time < 0 ? now_seconds() : wall_time, step, nullptr, &summary));
  {
    std::lock_guard<std::mutex> lk(w->mu);
    B200PS_REQUIRE(!w->stop, "events_add_scalars: writer is closed");
    w->pending += rec;
    ++w->records;
  }
  return 0;
}

// Block until everything added so far is in the file (and fflush'ed).
extern "C" int b200ps_events_flush(void* writer) {
  EventsWriter* w = static_cast<EventsWriter*>(writer);
  B200PS_REQUIRE(w, "events_flush: null writer");
  std::unique_lock<std::mutex> lk(w->mu);
  w->flush_now = true;
  w->cv.notify_all();
  w->cv.wait(lk, [&] { return w->pending.empty() && !w->flush_now && !w->writing; });
  return 0;
}

extern "C" int b200ps_events_close(void* writer) {
  EventsWriter* w = static_cast<EventsWriter*>(writer);
  if (!w) return 0;
  {
    std::lock_guard<std::mutex> lk(w->mu);
    w->stop = true;
  }
  w->cv.notify_all();
  if (w->worker.joinable()) w->worker.join();
  fclose(w->f);
  delete w;
  return 0;
}
