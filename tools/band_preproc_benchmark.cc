// tools/band_preproc_benchmark.cc -- C++ benchmark harness of the preprocessing path.
//
// The reference's tool (band/tool/band_benchmark.cc, benchmark.cc:78-100,102-276,530-563) is a JSON-configured CLI
// that drives the ENGINE with random tensors and prints average latency / FPS; it never touches band/buffer. This
// harness keeps its shape -- `band_preproc_benchmark <config.json>`, a report of avg latency and FPS -- for the
// band/buffer path: one host thread per GPU (frames are independent: no collective), a processor built through
// Band's C API (include/band_c_api.h), device-resident synthetic frames, CUDA-event timing on the processor's own
// stream, and a JSON report with the algorithmic HBM figure. It is a pure client of the two C ABIs: no kernel code,
// no CPU fallback.
//
// Config (all keys optional):
//   { "format": "NV12", "width": 1920, "height": 1080,
//     "operators": [["color_space_convert","RGB"], ["resize",224,224], ["normalize",127.5,127.5]],
//     "output": {"dtype": "float32", "dims": [1,224,224,3]},
//     "frames_per_step": 1024, "steps": 20, "warmup": 3, "devices": [0], "host_frames": false }
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <random>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "band_c_api.h"
#include "bandcu.h"

namespace {

// ---- a very small JSON reader (objects, arrays, strings, numbers, booleans): enough for the config above -------
struct Json {
  enum Kind { kNull, kNumber, kString, kBool, kArray, kObject } kind = kNull;
  double number = 0;
  bool boolean = false;
  std::string text;
  std::vector<Json> items;
  std::map<std::string, Json> fields;
  const Json* Get(const std::string& k) const {
    auto it = fields.find(k);
    return it == fields.end() ? nullptr : &it->second;
  }
};

struct JsonReader {
  const std::string& s;
  size_t i = 0;
  bool ok = true;
  explicit JsonReader(const std::string& str) : s(str) {}
  void Skip() { while (i < s.size() && (s[i] == ' ' || s[i] == '\n' || s[i] == '\t' || s[i] == '\r')) ++i; }
  bool Eat(char c) { Skip(); if (i < s.size() && s[i] == c) { ++i; return true; } return false; }
  Json Value() {
    Json v;
    Skip();
    if (i >= s.size()) { ok = false; return v; }
    const char c = s[i];
    if (c == '{') {
      ++i; v.kind = Json::kObject;
      if (Eat('}')) return v;
      do {
        Json k = Value();
        if (k.kind != Json::kString || !Eat(':')) { ok = false; return v; }
        v.fields[k.text] = Value();
      } while (ok && Eat(','));
      if (!Eat('}')) ok = false;
    } else if (c == '[') {
      ++i; v.kind = Json::kArray;
      if (Eat(']')) return v;
      do { v.items.push_back(Value()); } while (ok && Eat(','));
      if (!Eat(']')) ok = false;
    } else if (c == '"') {
      ++i; v.kind = Json::kString;
      while (i < s.size() && s[i] != '"') v.text.push_back(s[i++]);
      if (i >= s.size()) ok = false; else ++i;
    } else if (!s.compare(i, 4, "true")) { v.kind = Json::kBool; v.boolean = true; i += 4; }
    else if (!s.compare(i, 5, "false")) { v.kind = Json::kBool; i += 5; }
    else if (!s.compare(i, 4, "null")) { i += 4; }
    else {
      char* end = nullptr;
      v.number = strtod(s.c_str() + i, &end);
      if (end == s.c_str() + i) { ok = false; return v; }
      v.kind = Json::kNumber;
      i = static_cast<size_t>(end - s.c_str());
    }
    return v;
  }
};

// ---- configuration ------------------------------------------------------------------------------------------
struct OpSpec {
  BandImageProcessorBuilderField field;
  int i[4] = {0, 0, 0, 0};
  double f[2] = {0, 0};
};
struct Config {
  BandBufferFormat format = kBandNV12;
  int width = 1920, height = 1080;
  std::vector<OpSpec> ops;
  BandDataType out_type = kBandFloat32;
  std::vector<int> out_dims = {1, 224, 224, 3};
  int frames_per_step = 1024, steps = 20, warmup = 3;
  std::vector<int> devices = {0};
  bool host_frames = false;  // true: pinned host frames and tensors (the e2e path), false: HBM-resident
};

bool ParseFormat(const std::string& s, BandBufferFormat* f) {
  static const std::map<std::string, BandBufferFormat> m = {{"GRAY", kBandGrayScale}, {"RGB", kBandRGB}, {"RGBA", kBandRGBA},
                                                            {"YV12", kBandYV12}, {"YV21", kBandYV21}, {"NV21", kBandNV21},
                                                            {"NV12", kBandNV12}};
  auto it = m.find(s);
  if (it == m.end()) return false;
  *f = it->second;
  return true;
}

size_t FrameBytes(BandBufferFormat f, int w, int h) {
  const size_t cw = (w + 1) / 2, ch = (h + 1) / 2;
  switch (f) {
    case kBandGrayScale: return static_cast<size_t>(w) * h;
    case kBandRGB: return static_cast<size_t>(w) * h * 3;
    case kBandRGBA: return static_cast<size_t>(w) * h * 4;
    default: return static_cast<size_t>(w) * h + 2 * cw * ch;  // every 4:2:0 layout (buffer.cc:24-78)
  }
}

bool LoadConfig(const char* path, Config* c, std::string* err) {
  std::ifstream in(path);
  if (!in) { *err = std::string("cannot open ") + path; return false; }
  std::stringstream ss;
  ss << in.rdbuf();
  const std::string text = ss.str();
  JsonReader r(text);
  const Json root = r.Value();
  if (!r.ok || root.kind != Json::kObject) { *err = "config is not a JSON object"; return false; }
  if (const Json* v = root.Get("format")) if (!ParseFormat(v->text, &c->format)) { *err = "unknown format " + v->text; return false; }
  if (const Json* v = root.Get("width")) c->width = static_cast<int>(v->number);
  if (const Json* v = root.Get("height")) c->height = static_cast<int>(v->number);
  if (const Json* v = root.Get("frames_per_step")) c->frames_per_step = static_cast<int>(v->number);
  if (const Json* v = root.Get("steps")) c->steps = static_cast<int>(v->number);
  if (const Json* v = root.Get("warmup")) c->warmup = static_cast<int>(v->number);
  if (const Json* v = root.Get("host_frames")) c->host_frames = v->boolean;
  if (const Json* v = root.Get("devices")) {
    c->devices.clear();
    for (const Json& d : v->items) c->devices.push_back(static_cast<int>(d.number));
  }
  if (const Json* v = root.Get("output")) {
    if (const Json* t = v->Get("dtype")) {
      if (t->text == "float32") c->out_type = kBandFloat32;
      else if (t->text == "uint8") c->out_type = kBandUInt8;
      else { *err = "output.dtype must be float32 or uint8"; return false; }
    }
    if (const Json* d = v->Get("dims")) {
      c->out_dims.clear();
      for (const Json& x : d->items) c->out_dims.push_back(static_cast<int>(x.number));
    }
  }
  if (const Json* v = root.Get("operators")) {
    for (const Json& o : v->items) {
      if (o.kind != Json::kArray || o.items.empty()) { *err = "operator entries are arrays: [name, args...]"; return false; }
      const std::string& name = o.items[0].text;
      OpSpec op;
      const size_t argc = o.items.size() - 1;
      auto num = [&](size_t k) { return o.items[1 + k].number; };
      if (name == "crop" && argc == 4) { op.field = BAND_CROP; for (int k = 0; k < 4; ++k) op.i[k] = static_cast<int>(num(k)); }
      else if (name == "resize" && argc == 2) { op.field = BAND_RESIZE; op.i[0] = static_cast<int>(num(0)); op.i[1] = static_cast<int>(num(1)); }
      else if (name == "rotate" && argc == 1) { op.field = BAND_ROTATE; op.i[0] = static_cast<int>(num(0)); }
      else if (name == "flip" && argc == 2) { op.field = BAND_FLIP; op.i[0] = static_cast<int>(num(0)); op.i[1] = static_cast<int>(num(1)); }
      else if (name == "color_space_convert" && argc == 1) {
        BandBufferFormat f;
        if (!ParseFormat(o.items[1].text, &f)) { *err = "unknown format " + o.items[1].text; return false; }
        op.field = BAND_COLOR_SPACE_CONVERT; op.i[0] = static_cast<int>(f);
      }
      else if (name == "normalize" && argc == 2) { op.field = BAND_NORMALIZE; op.f[0] = num(0); op.f[1] = num(1); }
      else if (name == "data_type_convert" && argc == 0) { op.field = BAND_DATA_TYPE_CONVERT; }
      else if (name == "auto_orient" && argc == 0) { op.field = BAND_AUTO_ORIENT; }
      else { *err = "unknown operator or wrong argument count: " + name; return false; }
      c->ops.push_back(op);
    }
  } else {  // the north-star chain
    OpSpec a; a.field = BAND_COLOR_SPACE_CONVERT; a.i[0] = kBandRGB;
    OpSpec b; b.field = BAND_RESIZE; b.i[0] = 224; b.i[1] = 224;
    OpSpec n; n.field = BAND_NORMALIZE; n.f[0] = 127.5; n.f[1] = 127.5;
    c->ops = {a, b, n};
  }
  return true;
}

BandImageProcessor* BuildProcessor(const Config& c) {
  BandImageProcessorBuilder* b = BandImageProcessorBuilderCreate();
  BandStatus st = kBandOk;
  for (const OpSpec& op : c.ops) {
    switch (op.field) {
      case BAND_CROP: st = BandAddOperator(b, op.field, 4, op.i[0], op.i[1], op.i[2], op.i[3]); break;
      case BAND_RESIZE: case BAND_FLIP: st = BandAddOperator(b, op.field, 2, op.i[0], op.i[1]); break;
      case BAND_ROTATE: case BAND_COLOR_SPACE_CONVERT: st = BandAddOperator(b, op.field, 1, op.i[0]); break;
      case BAND_NORMALIZE: st = BandAddOperator(b, op.field, 2, op.f[0], op.f[1]); break;
      default: st = BandAddOperator(b, op.field, 0); break;
    }
    if (st != kBandOk) break;
  }
  BandImageProcessor* p = st == kBandOk ? BandImageProcessorBuilderBuild(b) : nullptr;
  BandImageProcessorBuilderDelete(b);
  return p;
}

// ---- one GPU, one host thread -----------------------------------------------------------------------------
struct DeviceResult {
  int device = 0;
  bool ok = false;
  std::string error;
  double ms_per_step = 0;      // CUDA events on the processor's stream (resident) / host clock (host_frames)
  long long launches = 0;
  int fused = 0;
};

void RunDevice(const Config& c, int device, DeviceResult* r) {
  r->device = device;
  auto fail = [&](const std::string& what) { r->error = what + ": " + BandGetLastError(); };
  if (BandCudaSetDevice(device) != kBandOk) return fail("BandCudaSetDevice");
  const size_t fb = FrameBytes(c.format, c.width, c.height);
  const size_t pitch = (fb + 255) / 256 * 256;  // 256-byte aligned frames: every source row is 16-byte aligned
  const int n = c.frames_per_step;
  void* frames = c.host_frames ? BandHostAlloc(pitch * n) : BandDeviceAlloc(pitch * n);
  if (!frames) return fail("frame allocation");
  {  // seeded synthetic frames: 8 distinct frames, repeated
    std::vector<unsigned char> uniq(8 * fb);
    std::mt19937 rng(0x42414E44u + static_cast<unsigned>(device));
    for (size_t k = 0; k < uniq.size(); k += 4) {
      const uint32_t v = rng();
      memcpy(&uniq[k], &v, uniq.size() - k < 4 ? uniq.size() - k : 4);
    }
    for (int f = 0; f < n; ++f) {
      unsigned char* dst = static_cast<unsigned char*>(frames) + pitch * f;
      if (c.host_frames) memcpy(dst, &uniq[(f % 8) * fb], fb);
      else if (BandCopyToDevice(dst, &uniq[(f % 8) * fb], fb) != kBandOk) return fail("BandCopyToDevice");
    }
  }
  BandImageProcessor* proc = BuildProcessor(c);
  if (!proc) return fail("BandImageProcessorBuilderBuild");
  std::vector<BandBuffer*> bufs(n);
  std::vector<BandTensor*> tens(n);
  for (int f = 0; f < n; ++f) {
    bufs[f] = BandBufferCreate();
    unsigned char* p = static_cast<unsigned char*>(frames) + pitch * f;
    const BandStatus st = c.host_frames ? BandBufferSetFromRawData(bufs[f], p, c.width, c.height, c.format)
                                        : BandBufferSetFromDeviceRawData(bufs[f], p, c.width, c.height, c.format);
    tens[f] = BandTensorCreate(c.out_type, c.out_dims.size(), c.out_dims.data(),
                               c.host_frames ? kBandHostMemory : kBandDeviceMemory);
    if (st != kBandOk || !tens[f]) return fail("buffer / tensor setup");
  }
  for (int k = 0; k < (c.warmup < 1 ? 1 : c.warmup); ++k)
    if (BandImageProcessorProcessBatch(proc, bufs.data(), tens.data(), n) != kBandOk) return fail("ProcessBatch");
  r->fused = BandImageProcessorIsFused(proc);
  const long long l0 = BandCudaLaunchCount();
  if (c.host_frames) {
    const auto t0 = std::chrono::steady_clock::now();
    for (int k = 0; k < c.steps; ++k)
      if (BandImageProcessorProcessBatch(proc, bufs.data(), tens.data(), n) != kBandOk) return fail("ProcessBatch");
    r->ms_per_step = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() / c.steps;
  } else {
    void* stream = BandImageProcessorGetStream(proc);
    void *e0 = nullptr, *e1 = nullptr;
    if (bandcu_event_create(&e0) != BANDCU_OK || bandcu_event_create(&e1) != BANDCU_OK) return fail("event");
    bandcu_event_record(e0, stream);
    for (int k = 0; k < c.steps; ++k)
      if (BandImageProcessorProcessBatchAsync(proc, bufs.data(), tens.data(), n) != kBandOk) return fail("ProcessBatchAsync");
    bandcu_event_record(e1, stream);
    if (BandImageProcessorSynchronize(proc) != kBandOk) return fail("Synchronize");
    float ms = 0;
    bandcu_event_elapsed_ms(e0, e1, &ms);
    r->ms_per_step = ms / c.steps;
    bandcu_event_destroy(e0);
    bandcu_event_destroy(e1);
  }
  r->launches = BandCudaLaunchCount() - l0;
  for (int f = 0; f < n; ++f) { BandBufferDelete(bufs[f]); BandTensorDelete(tens[f]); }
  BandImageProcessorDelete(proc);
  if (c.host_frames) BandHostFree(frames); else BandDeviceFree(frames);
  r->ok = true;
}

}  // namespace

int main(int argc, char** argv) {
  Config c;
  std::string err;
  if (argc >= 2 && !LoadConfig(argv[1], &c, &err)) {
    fprintf(stderr, "band_preproc_benchmark: %s\n", err.c_str());
    return 2;
  }
  if (c.ops.empty()) {
    OpSpec a; a.field = BAND_COLOR_SPACE_CONVERT; a.i[0] = kBandRGB;
    OpSpec b; b.field = BAND_RESIZE; b.i[0] = 224; b.i[1] = 224;
    OpSpec n; n.field = BAND_NORMALIZE; n.f[0] = 127.5; n.f[1] = 127.5;
    c.ops = {a, b, n};
  }
  std::vector<DeviceResult> results(c.devices.size());
  std::vector<std::thread> threads;
  const auto t0 = std::chrono::steady_clock::now();
  for (size_t k = 0; k < c.devices.size(); ++k) threads.emplace_back(RunDevice, std::cref(c), c.devices[k], &results[k]);
  for (std::thread& t : threads) t.join();
  const double wall_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  double slowest = 0;
  for (const DeviceResult& r : results) {
    if (!r.ok) { fprintf(stderr, "band_preproc_benchmark: device %d failed: %s\n", r.device, r.error.c_str()); return 1; }
    if (r.ms_per_step > slowest) slowest = r.ms_per_step;
  }
  size_t out_bytes = c.out_type == kBandFloat32 ? 4 : 1;
  for (int d : c.out_dims) out_bytes *= static_cast<size_t>(d);
  const double alg_bytes = static_cast<double>(FrameBytes(c.format, c.width, c.height) + out_bytes);
  const double fps = static_cast<double>(c.frames_per_step) * c.devices.size() / (slowest * 1e-3);
  printf("{\"tool\": \"band_preproc_benchmark\", \"n_gpus\": %zu, \"frames_per_step_per_gpu\": %d, \"steps\": %d, "
         "\"avg_latency_ms_per_step\": %.6f, \"frames_per_s\": %.1f, \"bytes_per_frame_algorithmic\": %.0f, "
         "\"achieved_GBps_per_gpu\": %.1f, \"host_frames\": %s, \"fused_single_launch\": %d, \"launches_per_step\": %.2f, "
         "\"timing\": \"%s, slowest device\", \"wall_s_total\": %.2f}\n",
         c.devices.size(), c.frames_per_step, c.steps, slowest, fps, alg_bytes,
         fps / c.devices.size() * alg_bytes / 1e9, c.host_frames ? "true" : "false", results[0].fused,
         static_cast<double>(results[0].launches) / c.steps,
         c.host_frames ? "host clock around synchronous ProcessBatch" : "CUDA events on the processor's stream", wall_s);
  return 0;
}
