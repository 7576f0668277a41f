// mof_multi.cu -- several GPUs behind one handle (include/mof.h, "mof_multi_*"): one mof_ctx and one
// persistent host worker thread per device, inside this library.  The hot path shards trivially
// (SURVEY.md 8e): every (stream, frame pair, section) is independent and the only state is each
// stream's previous frame, so there is no collective and no peer access -- a bag of pairs is cut into
// contiguous slices, camera stream s lives on device s % G, and every device copies its results D2H
// straight into its own disjoint slice of the caller's array.
//
// Replaces, for a multi-GPU box, what the reference does on one device per nodelet instance: the
// host<->device boundary of src/FftMethod.cpp:1415-1416 (upload) and :407 (result read-back).
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/mof.h"

namespace {

struct Worker {
  mof_ctx* ctx = nullptr;
  std::thread th;
  std::mutex mu;
  std::condition_variable cv;
  std::function<int()> job;
  bool has_job = false, done = true, quit = false;
  int rc = MOF_OK;

  void loop() {
    std::unique_lock<std::mutex> lk(mu);
    for (;;) {
      cv.wait(lk, [&] { return has_job || quit; });
      if (quit) return;
      std::function<int()> j = std::move(job);
      has_job = false;
      lk.unlock();
      const int r = j();
      lk.lock();
      rc = r;
      done = true;
      cv.notify_all();
    }
  }
  void post(std::function<int()> j) {
    std::lock_guard<std::mutex> lk(mu);
    job = std::move(j);
    has_job = true;
    done = false;
    cv.notify_all();
  }
  int wait() {
    std::unique_lock<std::mutex> lk(mu);
    cv.wait(lk, [&] { return done; });
    return rc;
  }
};

thread_local std::string g_multi_create_error;

}  // namespace

struct mof_multi {
  std::vector<Worker*> w;
  int F = 0, N = 0, S = 0;
  int n_streams = 0;
  std::string err;
  std::mutex call_mu;   // one call at a time, like a single context
};

namespace {

int mfail(mof_multi* m, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (m) m->err = buf;
  else g_multi_create_error = buf;
  return code;
}

// run one job per device (jobs[g] may be empty) and collect the first error
int run_all(mof_multi* m, std::vector<std::function<int()>>& jobs) {
  const int G = (int)m->w.size();
  for (int g = 0; g < G; g++)
    if (jobs[g]) m->w[g]->post(std::move(jobs[g]));
    else { std::lock_guard<std::mutex> lk(m->w[g]->mu); m->w[g]->rc = MOF_OK; m->w[g]->done = true; }
  int rc = MOF_OK;
  for (int g = 0; g < G; g++) {
    const int r = m->w[g]->wait();
    if (r != MOF_OK && rc == MOF_OK) {
      rc = r;
      m->err = std::string("device slot ") + std::to_string(g) + ": " + mof_last_error(m->w[g]->ctx);
    }
  }
  return rc;
}

// streams of one call grouped by owning device: positions k with ids[k] % G == g, local ids ids[k] / G
struct StreamSplit {
  std::vector<std::vector<int>> pos, local;
  // uniform stride between the positions of a device (k0, k0 + dk, ...), 0 when irregular
  std::vector<int> k0, dk;
};
StreamSplit split_streams(int G, int n, const int* ids) {
  StreamSplit sp;
  sp.pos.resize(G); sp.local.resize(G); sp.k0.assign(G, 0); sp.dk.assign(G, 0);
  for (int k = 0; k < n; k++) {
    const int g = ((ids[k] % G) + G) % G;
    sp.pos[g].push_back(k);
    sp.local[g].push_back(ids[k] / G);
  }
  for (int g = 0; g < G; g++) {
    const std::vector<int>& p = sp.pos[g];
    if (p.empty()) continue;
    sp.k0[g] = p[0];
    int d = p.size() > 1 ? p[1] - p[0] : 1;
    for (size_t i = 1; i < p.size(); i++)
      if (p[i] - p[i - 1] != d) { d = 0; break; }
    sp.dk[g] = d;
  }
  return sp;
}

}  // namespace

extern "C" {

int mof_multi_create(mof_multi** out, const mof_config* cfg, const int* devices, int n_devices) {
  if (!out || !cfg || !devices || n_devices < 1)
    return mfail(nullptr, MOF_ERR_INVALID, "mof_multi_create: null argument or no device");
  *out = nullptr;
  mof_multi* m = new mof_multi();
  m->n_streams = cfg->n_streams;
  for (int g = 0; g < n_devices; g++) {
    mof_config c = *cfg;
    c.device = devices[g];
    // stream s lives on device s % G as local stream s / G
    c.n_streams = cfg->n_streams > 0 ? (cfg->n_streams - g + n_devices - 1) / n_devices : 0;
    if (cfg->n_streams > 0 && c.n_streams == 0) c.n_streams = 1;
    Worker* w = new Worker();
    const int rc = mof_create(&w->ctx, &c);
    if (rc != MOF_OK) {
      const int code = mfail(nullptr, rc, "device %d: %s", devices[g], mof_last_error(nullptr));
      delete w;
      mof_multi_destroy(m);
      return code;
    }
    w->th = std::thread([w] { w->loop(); });
    m->w.push_back(w);
  }
  mof_get_geometry(m->w[0]->ctx, &m->F, &m->N, &m->S);
  *out = m;
  return MOF_OK;
}

void mof_multi_destroy(mof_multi* m) {
  if (!m) return;
  for (Worker* w : m->w) {
    {
      std::lock_guard<std::mutex> lk(w->mu);
      w->quit = true;
      w->cv.notify_all();
    }
    if (w->th.joinable()) w->th.join();
    mof_destroy(w->ctx);
    delete w;
  }
  delete m;
}

int mof_multi_device_count(const mof_multi* m) { return m ? (int)m->w.size() : 0; }

const char* mof_multi_last_error(const mof_multi* m) { return m ? m->err.c_str() : g_multi_create_error.c_str(); }

int mof_multi_get_geometry(const mof_multi* m, int* F, int* N, int* S) {
  if (!m) return MOF_ERR_INVALID;
  if (F) *F = m->F;
  if (N) *N = m->N;
  if (S) *S = m->S;
  return MOF_OK;
}

long long mof_multi_kernel_launches(const mof_multi* m) {
  long long n = 0;
  if (m)
    for (Worker* w : m->w) n += mof_kernel_launches(w->ctx);
  return n;
}

int mof_multi_fft_pairs(mof_multi* m, const uint8_t* curr, const uint8_t* prev, size_t pair_stride, int row_stride,
                        int n_pairs, double* shifts) {
  if (!m) return MOF_ERR_INVALID;
  std::unique_lock<std::mutex> call(m->call_mu, std::try_to_lock);
  if (!call.owns_lock()) return MOF_ERR_BUSY;
  if (n_pairs < 0 || (n_pairs > 0 && (!curr || !prev || !shifts))) return mfail(m, MOF_ERR_INVALID, "mof_multi_fft_pairs: bad argument");
  const int G = (int)m->w.size();
  std::vector<std::function<int()>> jobs(G);
  const int per = (n_pairs + G - 1) / G;   // contiguous slices
  for (int g = 0; g < G; g++) {
    const int p0 = g * per, n = n_pairs - p0 < per ? n_pairs - p0 : per;
    if (n <= 0) continue;
    mof_ctx* ctx = m->w[g]->ctx;
    const uint8_t* c = curr + (size_t)p0 * pair_stride;
    const uint8_t* p = prev + (size_t)p0 * pair_stride;
    double* o = shifts + (size_t)p0 * m->S * 2;
    jobs[g] = [=] { return mof_fft_pairs(ctx, c, p, pair_stride, row_stride, n, o); };
  }
  return run_all(m, jobs);
}

int mof_multi_fft_stream_push(mof_multi* m, int n, const int* ids, const uint8_t* frames, size_t frame_stride,
                              int row_stride, double* shifts) {
  return mof_multi_fft_stream_push_many(m, n, ids, 1, frames, frame_stride, frame_stride, row_stride, shifts);
}

int mof_multi_fft_stream_push_many(mof_multi* m, int n, const int* ids, int T, const uint8_t* frames,
                                   size_t stream_stride, size_t frame_stride, int row_stride, double* shifts) {
  if (!m) return MOF_ERR_INVALID;
  std::unique_lock<std::mutex> call(m->call_mu, std::try_to_lock);
  if (!call.owns_lock()) return MOF_ERR_BUSY;
  if (n < 0 || T < 1 || (n > 0 && (!ids || !frames || !shifts)))
    return mfail(m, MOF_ERR_INVALID, "mof_multi_fft_stream_push_many: bad argument");
  for (int k = 0; k < n; k++)
    if (ids[k] < 0 || ids[k] >= m->n_streams) return mfail(m, MOF_ERR_INVALID, "stream id %d out of range", ids[k]);
  const int G = (int)m->w.size(), S = m->S;
  StreamSplit sp = split_streams(G, n, ids);
  std::vector<std::function<int()>> jobs(G);
  const size_t per_stream = (size_t)T * S * 2;   // doubles
  for (int g = 0; g < G; g++) {
    const int ng = (int)sp.pos[g].size();
    if (!ng) continue;
    mof_ctx* ctx = m->w[g]->ctx;
    const std::vector<int> pos = sp.pos[g], local = sp.local[g];
    const int k0 = sp.k0[g], dk = sp.dk[g];
    if (dk == 1) {
      // this device's streams are adjacent in the call: one call, results straight into the caller's slice
      jobs[g] = [=] {
        return mof_fft_stream_push_many(ctx, ng, local.data(), T, frames + (size_t)k0 * stream_stride, stream_stride,
                                        frame_stride, row_stride, shifts + (size_t)k0 * per_stream);
      };
    } else if (dk > 1) {
      // interleaved ids (0..n-1 over G devices): uniform stride between this device's streams; the results
      // are gathered in a device-ordered block and scattered per stream on the host (T*S*16 bytes each)
      jobs[g] = [=] {
        std::vector<double> tmp((size_t)ng * per_stream);
        int rc = mof_fft_stream_push_many(ctx, ng, local.data(), T, frames + (size_t)k0 * stream_stride,
                                          stream_stride * dk, frame_stride, row_stride, tmp.data());
        if (rc) return rc;
        for (int i = 0; i < ng; i++)
          std::copy(tmp.begin() + (size_t)i * per_stream, tmp.begin() + (size_t)(i + 1) * per_stream,
                    shifts + (size_t)pos[i] * per_stream);
        return (int)MOF_OK;
      };
    } else {
      // any other id pattern: one call per stream, results straight into the caller's slice of that stream
      jobs[g] = [=] {
        for (int i = 0; i < ng; i++) {
          int rc = mof_fft_stream_push_many(ctx, 1, &local[i], T, frames + (size_t)pos[i] * stream_stride, stream_stride,
                                            frame_stride, row_stride, shifts + (size_t)pos[i] * per_stream);
          if (rc) return rc;
        }
        return (int)MOF_OK;
      };
    }
  }
  return run_all(m, jobs);
}

int mof_multi_fft_stream_push_many_raw(mof_multi* m, int n, const int* ids, int T, const mof_frontend_desc* d,
                                       const uint8_t* raw, double* shifts) {
  if (!m) return MOF_ERR_INVALID;
  std::unique_lock<std::mutex> call(m->call_mu, std::try_to_lock);
  if (!call.owns_lock()) return MOF_ERR_BUSY;
  if (n < 0 || T < 1 || !d || (n > 0 && (!ids || !raw || !shifts)))
    return mfail(m, MOF_ERR_INVALID, "mof_multi_fft_stream_push_many_raw: bad argument");
  for (int k = 0; k < n; k++)
    if (ids[k] < 0 || ids[k] >= m->n_streams) return mfail(m, MOF_ERR_INVALID, "stream id %d out of range", ids[k]);
  const int G = (int)m->w.size(), S = m->S;
  StreamSplit sp = split_streams(G, n, ids);
  std::vector<std::function<int()>> jobs(G);
  const size_t per_stream = (size_t)T * S * 2;
  const mof_frontend_desc desc = *d;
  for (int g = 0; g < G; g++) {
    const int ng = (int)sp.pos[g].size();
    if (!ng) continue;
    mof_ctx* ctx = m->w[g]->ctx;
    const std::vector<int> pos = sp.pos[g], local = sp.local[g];
    const int k0 = sp.k0[g], dk = sp.dk[g];
    jobs[g] = [=] {
      if (dk == 1)
        return mof_fft_stream_push_many_raw(ctx, ng, local.data(), T, &desc, raw + (size_t)k0 * T * desc.image_stride,
                                            shifts + (size_t)k0 * per_stream);
      for (int i = 0; i < ng; i++) {
        int rc = mof_fft_stream_push_many_raw(ctx, 1, &local[i], T, &desc, raw + (size_t)pos[i] * T * desc.image_stride,
                                              shifts + (size_t)pos[i] * per_stream);
        if (rc) return rc;
      }
      return (int)MOF_OK;
    };
  }
  return run_all(m, jobs);
}

}  // extern "C"
