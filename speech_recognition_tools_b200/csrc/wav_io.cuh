// Host-side WAV ingest of the CLI pipeline (no device code).
//
// Reference: src/featgen/computeFDLPSpectrogram.py:129-152 reads every utterance with scipy.io.wavfile.read.  For the
// plain files of a Kaldi wav.scp (mono RIFF, PCM16 or IEEE float32) the shim needs two things per file: where the
// samples are (header probe) and the samples themselves, packed back to back in pinned memory.  In Python the probes
// cost 56 us per file and cannot be spread over threads (the interpreter lock); here a few threads walk the list.
// Anything that is not such a file is reported as format -1 and goes through the reference's reader on the Python side.
#pragma once
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace fdlp {

inline uint32_t rd_u32(const unsigned char* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
inline uint32_t rd_u16(const unsigned char* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8); }

// info = {srate, format, data_off, n_samples}; format 0: int16, 1: float32, -1: not handled here
inline void wav_probe_one(const char* path, int64_t* info) {
  info[0] = 0; info[1] = -1; info[2] = 0; info[3] = 0;
  const int fd = open(path, O_RDONLY | O_CLOEXEC);
  if (fd < 0) return;
  unsigned char head[4096];
  struct stat st;
  const ssize_t got = pread(fd, head, sizeof(head), 0);
  const int st_rc = fstat(fd, &st);
  close(fd);
  if (got < 44 || st_rc != 0) return;
  if (memcmp(head, "RIFF", 4) != 0 || memcmp(head + 8, "WAVE", 4) != 0) return;
  int64_t pos = 12;
  bool have_fmt = false;
  uint32_t tag = 0, nch = 0, sr = 0, bits = 0;
  while (pos + 8 <= got) {
    const unsigned char* c = head + pos;
    const uint32_t csz = rd_u32(c + 4);
    if (memcmp(c, "fmt ", 4) == 0) {
      if (pos + 8 + 16 > got) return;
      tag = rd_u16(c + 8); nch = rd_u16(c + 10); sr = rd_u32(c + 12); bits = rd_u16(c + 22);
      have_fmt = true;
    } else if (memcmp(c, "data", 4) == 0) {
      if (!have_fmt || nch != 1) return;
      int fmt, isz;
      if (tag == 1 && bits == 16) { fmt = 0; isz = 2; }
      else if (tag == 3 && bits == 32) { fmt = 1; isz = 4; }
      else return;
      int64_t nbytes = (int64_t)st.st_size - (pos + 8);
      if ((int64_t)csz < nbytes) nbytes = csz;
      if (nbytes < 0) nbytes = 0;
      info[0] = sr; info[1] = fmt; info[2] = pos + 8; info[3] = nbytes / isz;
      return;
    }
    pos += 8 + (int64_t)csz + (csz & 1);
  }
}

template <typename F>
inline void run_indexed(int n, int threads, F&& f) {
  if (threads < 1) threads = 1;
  if (threads > n) threads = n;
  if (threads <= 1) {
    for (int i = 0; i < n; ++i) f(i);
    return;
  }
  std::atomic<int> next{0};
  std::vector<std::thread> pool;
  pool.reserve(threads);
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&]() {
      for (int i = next.fetch_add(1); i < n; i = next.fetch_add(1)) f(i);
    });
  for (auto& th : pool) th.join();
}

inline int wav_read_one(const char* path, int64_t off, int64_t nbytes, unsigned char* dst) {
  const int fd = open(path, O_RDONLY | O_CLOEXEC);
  if (fd < 0) return 1;
  int64_t got = 0;
  while (got < nbytes) {
    const ssize_t r = pread(fd, dst + got, (size_t)(nbytes - got), off + got);
    if (r <= 0) break;
    got += r;
  }
  close(fd);
  return got == nbytes ? 0 : 1;
}

}  // namespace fdlp
