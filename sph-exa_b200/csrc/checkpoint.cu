// Restart files: dump and reload of the conserved particle state.
//
// Reference: the main loop writes a restart dump through IFileWriter (main/src/sphexa/sphexa.cpp:232-246,
// main/src/io/ifile_io_hdf5.cpp): step attributes from ParticlesData::loadOrStoreAttributes
// (sph/include/sph/particles_data.hpp:142-193), Box::loadOrStore (domain/include/cstone/sfc/box.hpp:168-175),
// TurbulenceData::loadOrStore (sph/include/sph/hydro_turb/turbulence_data.hpp:90-125), then one dataset per conserved
// field of the propagator (ve_hydro.hpp:75, conservedFields()).  Restart (init/file_init.hpp:93-140) reads an equal
// contiguous slice of every dataset per rank and lets the first domain sync redistribute.
//
// H5Part/HDF5 do not exist on this box, so the container is a plain little-endian binary file with the same logical
// content and the same names:
//
//   char magic[8] "SPHB200R" | u32 version | u32 numBlocks
//   numBlocks x { char name[40] | u32 kind (0 attribute, 1 particle field) | u32 type (0 f64, 1 i64, 2 bytes, 3 f32)
//                 | u64 count | u64 payload offset from the start of the file }
//   payloads, each 8-byte aligned
//
// Host code only; device state moves through the public sim calls (download / upload).
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "common.cuh"

namespace
{

constexpr char kMagic[8]     = {'S', 'P', 'H', 'B', '2', '0', '0', 'R'};
constexpr uint32_t kVersion  = 1;
constexpr uint32_t kNameSize = 40;

enum BlockKind : uint32_t { B_ATTRIBUTE = 0, B_FIELD = 1 };
enum BlockType : uint32_t { T_F64 = 0, T_I64 = 1, T_BYTES = 2, T_F32 = 3 };

size_t typeBytes(uint32_t t) { return t == T_BYTES ? 1 : (t == T_F32 ? 4 : 8); }

struct BlockHeader
{
    char name[kNameSize];
    uint32_t kind, type;
    uint64_t count, offset;
};
static_assert(sizeof(BlockHeader) == kNameSize + 24, "directory entries are written as they are");

struct FileCloser
{
    void operator()(std::FILE* f) const
    {
        if (f) std::fclose(f);
    }
};
using FilePtr = std::unique_ptr<std::FILE, FileCloser>;

//! @brief names of the conserved fields in the order of sphb200_fields, with a flag: always f64 (x, y, z, temp)
struct FieldName
{
    const char* name;
    bool alwaysF64;
    void* sphb200_fields::*real;
    double* sphb200_fields::*dbl;
};

const FieldName kConserved[] = {{"x", true, nullptr, &sphb200_fields::x},
                                {"y", true, nullptr, &sphb200_fields::y},
                                {"z", true, nullptr, &sphb200_fields::z},
                                {"h", false, &sphb200_fields::h, nullptr},
                                {"m", false, &sphb200_fields::m, nullptr},
                                {"vx", false, &sphb200_fields::vx, nullptr},
                                {"vy", false, &sphb200_fields::vy, nullptr},
                                {"vz", false, &sphb200_fields::vz, nullptr},
                                {"x_m1", false, &sphb200_fields::x_m1, nullptr},
                                {"y_m1", false, &sphb200_fields::y_m1, nullptr},
                                {"z_m1", false, &sphb200_fields::z_m1, nullptr},
                                {"temp", true, nullptr, &sphb200_fields::temp},
                                {"du_m1", false, &sphb200_fields::du_m1, nullptr},
                                {"alpha", false, &sphb200_fields::alpha, nullptr}};
constexpr int kNumConservedFields = sizeof(kConserved) / sizeof(kConserved[0]);

void* hostPtr(const sphb200_fields& f, const FieldName& fn)
{
    return fn.alwaysF64 ? static_cast<void*>(f.*(fn.dbl)) : f.*(fn.real);
}

void setHostPtr(sphb200_fields& f, const FieldName& fn, void* p)
{
    if (fn.alwaysF64) f.*(fn.dbl) = static_cast<double*>(p);
    else f.*(fn.real) = p;
}

//! @brief collects blocks, then writes directory and payloads in one pass
class Writer
{
public:
    void add(const std::string& name, uint32_t kind, uint32_t type, uint64_t count, const void* data)
    {
        if (name.size() >= kNameSize) throw std::runtime_error("restart: name too long: " + name);
        BlockHeader h;
        std::memset(&h, 0, sizeof(h));
        std::memcpy(h.name, name.data(), name.size());
        h.kind  = kind;
        h.type  = type;
        h.count = count;
        headers_.push_back(h);
        data_.push_back(data);
    }

    void f64(const std::string& name, double v)
    {
        scalars_.push_back(std::make_unique<uint64_t>());
        std::memcpy(scalars_.back().get(), &v, 8);
        add(name, B_ATTRIBUTE, T_F64, 1, scalars_.back().get());
    }

    void i64(const std::string& name, int64_t v)
    {
        scalars_.push_back(std::make_unique<uint64_t>());
        std::memcpy(scalars_.back().get(), &v, 8);
        add(name, B_ATTRIBUTE, T_I64, 1, scalars_.back().get());
    }

    void write(const char* path)
    {
        // write to a temporary name and rename: a dump interrupted half-way never replaces a good file
        std::string tmp = std::string(path) + ".part";
        FilePtr f(std::fopen(tmp.c_str(), "wb"));
        if (!f) throw std::runtime_error(std::string("restart: cannot open for writing: ") + tmp);
        uint64_t pos = sizeof(kMagic) + 8 + headers_.size() * sizeof(BlockHeader);
        for (auto& h : headers_)
        {
            h.offset = pos;
            pos += (h.count * typeBytes(h.type) + 7) & ~uint64_t(7);
        }
        uint32_t nb = uint32_t(headers_.size());
        put(f.get(), kMagic, sizeof(kMagic));
        put(f.get(), &kVersion, 4);
        put(f.get(), &nb, 4);
        put(f.get(), headers_.data(), headers_.size() * sizeof(BlockHeader));
        const char zeros[8] = {};
        for (size_t b = 0; b < headers_.size(); ++b)
        {
            size_t bytes = headers_[b].count * typeBytes(headers_[b].type);
            put(f.get(), data_[b], bytes);
            put(f.get(), zeros, ((bytes + 7) & ~size_t(7)) - bytes);
        }
        if (std::fflush(f.get()) != 0) throw std::runtime_error("restart: write failed: " + tmp);
        f.reset();
        if (std::rename(tmp.c_str(), path) != 0) throw std::runtime_error(std::string("restart: cannot rename to ") + path);
    }

private:
    static void put(std::FILE* f, const void* p, size_t bytes)
    {
        if (bytes && std::fwrite(p, 1, bytes, f) != bytes) throw std::runtime_error("restart: write failed");
    }

    std::vector<BlockHeader> headers_;
    std::vector<const void*> data_;
    std::vector<std::unique_ptr<uint64_t>> scalars_;
};

} // namespace

struct sphb200_restart
{
    std::string path;
    FilePtr file;
    uint64_t fileSize = 0;
    std::vector<BlockHeader> blocks;

    const BlockHeader* find(const char* name, uint32_t kind) const
    {
        for (auto& b : blocks)
            if (b.kind == kind && std::strncmp(b.name, name, kNameSize) == 0) return &b;
        return nullptr;
    }

    void read(uint64_t offset, void* dst, size_t bytes) const
    {
        if (bytes == 0) return;
        if (fseeko(file.get(), off_t(offset), SEEK_SET) != 0 || std::fread(dst, 1, bytes, file.get()) != bytes)
            throw std::runtime_error("restart: short read in " + path);
    }

    bool f64(const char* name, double* v) const
    {
        const BlockHeader* b = find(name, B_ATTRIBUTE);
        if (!b || b->type != T_F64 || b->count < 1) return false;
        read(b->offset, v, 8);
        return true;
    }

    bool i64(const char* name, int64_t* v) const
    {
        const BlockHeader* b = find(name, B_ATTRIBUTE);
        if (!b || b->type != T_I64 || b->count < 1) return false;
        read(b->offset, v, 8);
        return true;
    }

    double requireF64(const char* name) const
    {
        double v;
        if (!f64(name, &v)) throw std::runtime_error("restart: attribute missing in " + path + ": " + name);
        return v;
    }

    int64_t requireI64(const char* name) const
    {
        int64_t v;
        if (!i64(name, &v)) throw std::runtime_error("restart: attribute missing in " + path + ": " + name);
        return v;
    }
};

using sphb200::setError;

#define RESTART_TRY try {
#define RESTART_CATCH                                                                                                 \
    return 0;                                                                                                         \
    }                                                                                                                 \
    catch (std::exception & e)                                                                                        \
    {                                                                                                                 \
        setError(e.what());                                                                                           \
        return 1;                                                                                                     \
    }

#define RESTART_RC(call)                                                                                              \
    do                                                                                                                \
    {                                                                                                                 \
        if ((call) != 0) throw std::runtime_error(sphb200_last_error());                                              \
    } while (0)

extern "C" int sphb200_restart_write_bdt(const char* path, const sphb200_restart_info* info,
                                         const sphb200_hydro_const* hcIn, const sphb200_turb* t, int rank,
                                         int numRanks, const sphb200_fields* host, const sphb200_restart_bdt* bdt)
{
    RESTART_TRY
    const sphb200_hydro_const& hc = *hcIn;
    size_t n  = info->numParticles;
    size_t eb = info->dtype == SPHB200_F64 ? 8 : 4;

    Writer w;
    // ParticlesData::loadOrStoreAttributes (particles_data.hpp:169-189), same names
    w.i64("iteration", int64_t(info->iteration));
    w.i64("numParticlesGlobal", int64_t(info->numParticlesGlobal));
    w.i64("ng0", hc.ng0);
    w.i64("ngmax", hc.ngmax);
    w.f64("time", info->time);
    w.f64("minDt", info->minDt);
    w.f64("minDt_m1", info->minDt_m1);
    w.f64("Kcour", hc.Kcour);
    w.f64("Krho", hc.Krho);
    w.f64("gravConstant", 0.0);
    w.f64("gamma", hc.gamma);
    w.f64("muiConst", hc.muiConst);
    w.f64("alphamin", hc.alphamin);
    w.f64("alphamax", hc.alphamax);
    w.f64("decay_constant", hc.decay_constant);
    // not file attributes in the reference (compile-time or derived there), needed to rebuild sphb200_hydro_const
    w.f64("kernelNormalization", hc.K);
    w.f64("Atmin", hc.Atmin);
    w.f64("Atmax", hc.Atmax);
    w.f64("ramp", hc.ramp);
    w.i64("avClean", hc.avClean);
    w.i64("precision", info->dtype);
    w.i64("rank", rank);
    w.i64("numRanks", numRanks);
    w.i64("numParticles", int64_t(n));
    // Box::loadOrStore (box.hpp:170-171)
    w.add("box", B_ATTRIBUTE, T_F64, 6, info->box.lim);
    char bc[3] = {char(info->box.bc[0]), char(info->box.bc[1]), char(info->box.bc[2])};
    w.add("boundaryType", B_ATTRIBUTE, T_BYTES, 3, bc);

    // TurbulenceData::loadOrStore (turbulence_data.hpp:92-115)
    std::vector<double> modes, amplitudes, phases;
    std::string rng;
    double scalars[4];
    if (t)
    {
        int nm = sphb200_turb_num_modes(t);
        modes.resize(3 * size_t(nm));
        amplitudes.resize(nm);
        phases.resize(6 * size_t(nm));
        RESTART_RC(sphb200_turb_get(t, scalars, modes.data(), amplitudes.data(), phases.data(), nullptr, nullptr));
        rng.resize(sphb200_turb_rng_state(t, nullptr, 0));
        sphb200_turb_rng_state(t, rng.data(), rng.size());
        w.f64("turbulence::variance", scalars[0]);
        w.f64("turbulence::decayTime", scalars[1]);
        w.f64("turbulence::solWeight", scalars[2]);
        w.f64("turbulence::solWeightNorm", scalars[3]);
        w.i64("turbulence::numModes", nm);
        w.add("turbulence::modes", B_ATTRIBUTE, T_F64, modes.size(), modes.data());
        w.add("turbulence::amplitudes", B_ATTRIBUTE, T_F64, amplitudes.size(), amplitudes.data());
        w.add("turbulence::phases", B_ATTRIBUTE, T_F64, phases.size(), phases.data());
        w.add("rngEngineState", B_ATTRIBUTE, T_BYTES, rng.size(), rng.data());
    }

    for (int k = 0; k < kNumConservedFields; ++k)
    {
        const void* src = hostPtr(*host, kConserved[k]);
        if (!src && n) throw std::runtime_error(std::string("restart: conserved field missing: ") + kConserved[k].name);
        w.add(kConserved[k].name, B_FIELD, (kConserved[k].alwaysF64 || eb == 8) ? T_F64 : T_F32, n, src);
    }
    if (bdt)
    {
        // Timestep::loadOrStore with the prefix "ts::" (timestep.h:29-33), and the conserved field "rung"
        if (!bdt->rung && n) throw std::runtime_error("restart: conserved field missing: rung");
        w.i64("ts::numRungs", bdt->numRungs);
        w.add("ts::dt_m1", B_ATTRIBUTE, T_F32, 4, bdt->dt_m1);
        w.add("rung", B_FIELD, T_BYTES, n, bdt->rung);
    }
    w.write(path);
    RESTART_CATCH
}

extern "C" int sphb200_restart_write(const char* path, const sphb200_restart_info* info,
                                     const sphb200_hydro_const* hc, const sphb200_turb* t, int rank, int numRanks,
                                     const sphb200_fields* host)
{
    return sphb200_restart_write_bdt(path, info, hc, t, rank, numRanks, host, nullptr);
}

extern "C" int sphb200_sim_save(sphb200_sim* s, const char* path, uint64_t iteration, uint64_t numParticlesGlobal)
{
    RESTART_TRY
    sphb200_sim_config cfg;
    sphb200_sim_get_config(s, &cfg);
    size_t n  = sphb200_sim_end_index(s) - sphb200_sim_start_index(s);
    size_t eb = cfg.dtype == SPHB200_F64 ? 8 : 4;

    sphb200_restart_info info;
    std::memset(&info, 0, sizeof(info));
    info.iteration          = iteration;
    info.numParticlesGlobal = numParticlesGlobal;
    info.numParticles       = n;
    info.dtype              = cfg.dtype;
    info.time               = sphb200_sim_ttot(s);
    info.minDt              = cfg.hc.minDt;
    info.minDt_m1           = cfg.hc.minDt_m1;
    sphb200_sim_get_box(s, &info.box);

    // conserved fields of the owned range, one device-to-host copy each
    std::vector<std::vector<char>> host(kNumConservedFields);
    sphb200_fields table;
    std::memset(&table, 0, sizeof(table));
    for (int k = 0; k < kNumConservedFields; ++k)
    {
        size_t bytes = kConserved[k].alwaysF64 ? 8 : eb;
        host[k].resize(std::max<size_t>(n * bytes, 8));
        setHostPtr(table, kConserved[k], host[k].data());
    }
    RESTART_RC(sphb200_sim_download(s, &table, nullptr));

    // HydroVeBdtProp::save (ve_hydro_bdt.hpp:153)
    sphb200_bdt_info bi;
    RESTART_RC(sphb200_sim_bdt_info(s, &bi));
    sphb200_restart_bdt bdt;
    std::vector<uint8_t> rung;
    bool haveBdt = sphb200_sim_bdt_rungs(s) != nullptr;
    // Only ts::numRungs, ts::dt_m1 and the rung field travel (like HydroVeBdtProp::save): substep, elapsed time, drift
    // times and rung ranges do not, and the higher rungs sit at drifted positions in the middle of a hierarchy.  The
    // reference only writes output when Propagator::isSynced() holds (sphexa.cpp:153-181); a dump taken anywhere
    // else would resume into a different run without any error, so it is refused.
    if (haveBdt && !bi.isSynced)
        throw std::runtime_error("restart: the block-time-step hierarchy is at substep " + std::to_string(bi.substep) +
                                 " of " + std::to_string(1 << std::max(bi.numRungs - 1, 0)) +
                                 "; dumps are only taken where the propagator is synced (isSynced)");
    if (haveBdt)
    {
        rung.resize(std::max<size_t>(n, 1));
        RESTART_RC(sphb200_sim_bdt_download_rungs(s, rung.data()));
        bdt.numRungs = bi.numRungs;
        std::copy(bi.dt_m1, bi.dt_m1 + 4, bdt.dt_m1);
        bdt.rung = rung.data();
    }
    RESTART_RC(sphb200_restart_write_bdt(path, &info, &cfg.hc, sphb200_sim_turbulence(s), cfg.rank, cfg.numRanks,
                                         &table, haveBdt ? &bdt : nullptr));
    RESTART_CATCH
}

extern "C" int sphb200_restart_open(const char* path, sphb200_restart** out)
{
    RESTART_TRY
    auto r  = std::make_unique<sphb200_restart>();
    r->path = path;
    r->file.reset(std::fopen(path, "rb"));
    if (!r->file) throw std::runtime_error(std::string("restart: cannot open ") + path);
    fseeko(r->file.get(), 0, SEEK_END);
    r->fileSize = uint64_t(ftello(r->file.get()));
    char magic[8];
    uint32_t version = 0, nb = 0;
    if (r->fileSize < 16) throw std::runtime_error(std::string("restart: not a restart file: ") + path);
    r->read(0, magic, 8);
    r->read(8, &version, 4);
    r->read(12, &nb, 4);
    if (std::memcmp(magic, kMagic, 8) != 0) throw std::runtime_error(std::string("restart: not a restart file: ") + path);
    if (version != kVersion) throw std::runtime_error("restart: unsupported file version " + std::to_string(version));
    if (16 + uint64_t(nb) * sizeof(BlockHeader) > r->fileSize) throw std::runtime_error("restart: truncated directory");
    r->blocks.resize(nb);
    r->read(16, r->blocks.data(), nb * sizeof(BlockHeader));
    for (auto& b : r->blocks)
    {
        b.name[kNameSize - 1] = 0;
        // no sums or products of untrusted sizes: they would wrap for a corrupt directory entry
        if (b.type > T_F32 || b.offset > r->fileSize || b.count > (r->fileSize - b.offset) / typeBytes(b.type))
            throw std::runtime_error(std::string("restart: truncated or corrupt block: ") + b.name);
    }
    *out = r.release();
    RESTART_CATCH
}

extern "C" void sphb200_restart_close(sphb200_restart* r) { delete r; }

extern "C" int sphb200_restart_get_info(const sphb200_restart* r, sphb200_restart_info* out)
{
    RESTART_TRY
    std::memset(out, 0, sizeof(*out));
    out->iteration          = uint64_t(r->requireI64("iteration"));
    out->numParticlesGlobal = uint64_t(r->requireI64("numParticlesGlobal"));
    out->numParticles       = uint64_t(r->requireI64("numParticles"));
    out->dtype              = int(r->requireI64("precision"));
    out->hasTurbulence      = r->find("turbulence::phases", B_ATTRIBUTE) != nullptr;
    out->time               = r->requireF64("time");
    out->minDt              = r->requireF64("minDt");
    out->minDt_m1           = r->requireF64("minDt_m1");
    const BlockHeader* b    = r->find("box", B_ATTRIBUTE);
    const BlockHeader* t    = r->find("boundaryType", B_ATTRIBUTE);
    if (!b || b->count != 6 || !t || t->count != 3) throw std::runtime_error("restart: box attributes missing");
    double lim[6];
    char bc[3];
    r->read(b->offset, lim, sizeof(lim));
    r->read(t->offset, bc, 3);
    int bci[3] = {bc[0], bc[1], bc[2]};
    sphb200_make_box(lim, bci, &out->box);
    RESTART_CATCH
}

extern "C" int64_t sphb200_restart_attribute(const sphb200_restart* r, const char* name, void* out, size_t capacity)
{
    try
    {
        const BlockHeader* b = r->find(name, B_ATTRIBUTE);
        if (!b) return -1;
        if (out) r->read(b->offset, out, std::min<uint64_t>(capacity, b->count) * typeBytes(b->type));
        return int64_t(b->count);
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return -2;
    }
}

extern "C" int sphb200_restart_read_fields(const sphb200_restart* r, uint64_t first, uint64_t count, int dtype,
                                           const sphb200_fields* host)
{
    RESTART_TRY
    std::vector<char> staging;
    for (int k = 0; k < kNumConservedFields; ++k)
    {
        void* dst = hostPtr(*host, kConserved[k]);
        if (!dst) continue;
        const BlockHeader* b = r->find(kConserved[k].name, B_FIELD);
        if (!b) throw std::runtime_error(std::string("restart: field missing in ") + r->path + ": " + kConserved[k].name);
        if (first > b->count || count > b->count - first)
            throw std::runtime_error(std::string("restart: range beyond the end of field ") + b->name);
        size_t want = (kConserved[k].alwaysF64 || dtype == SPHB200_F64) ? 8 : 4;
        size_t have = typeBytes(b->type);
        if (want == have) { r->read(b->offset + first * have, dst, count * have); }
        else
        {
            // the file holds the other precision: convert element-wise, like fileutils::readField
            staging.resize(count * have);
            r->read(b->offset + first * have, staging.data(), count * have);
            if (want == 8)
            {
                auto* src = reinterpret_cast<const float*>(staging.data());
                auto* d   = static_cast<double*>(dst);
                for (uint64_t i = 0; i < count; ++i)
                    d[i] = double(src[i]);
            }
            else
            {
                auto* src = reinterpret_cast<const double*>(staging.data());
                auto* d   = static_cast<float*>(dst);
                for (uint64_t i = 0; i < count; ++i)
                    d[i] = float(src[i]);
            }
        }
    }
    RESTART_CATCH
}

extern "C" int sphb200_restart_read_bytes(const sphb200_restart* r, const char* name, uint64_t first, uint64_t count,
                                          uint8_t* out)
{
    RESTART_TRY
    const BlockHeader* b = r->find(name, B_FIELD);
    if (!b || b->type != T_BYTES) throw std::runtime_error(std::string("restart: field missing in ") + r->path + ": " + name);
    if (first > b->count || count > b->count - first)
        throw std::runtime_error(std::string("restart: range beyond the end of field ") + b->name);
    r->read(b->offset + first, out, count);
    RESTART_CATCH
}

extern "C" int sphb200_restart_apply(const sphb200_restart* r, sphb200_sim* s)
{
    RESTART_TRY
    sphb200_restart_info info;
    RESTART_RC(sphb200_restart_get_info(r, &info));
    sphb200_sim_config cfg;
    sphb200_sim_get_config(s, &cfg);
    sphb200_hydro_const hc = cfg.hc;
    // optional attributes keep the value the sim was created with when absent (optionalIO, particles_data.hpp:145-167)
    int64_t iv;
    if (r->i64("ng0", &iv)) hc.ng0 = unsigned(iv);
    if (r->i64("ngmax", &iv)) hc.ngmax = unsigned(iv);
    r->f64("Kcour", &hc.Kcour);
    r->f64("Krho", &hc.Krho);
    r->f64("gamma", &hc.gamma);
    r->f64("muiConst", &hc.muiConst);
    r->f64("alphamin", &hc.alphamin);
    r->f64("alphamax", &hc.alphamax);
    r->f64("decay_constant", &hc.decay_constant);
    r->f64("Atmin", &hc.Atmin);
    r->f64("Atmax", &hc.Atmax);
    r->f64("ramp", &hc.ramp);
    r->f64("kernelNormalization", &hc.K);
    // avClean selects other kernels and another field set: fixed when the simulation is created, must agree
    if (r->i64("avClean", &iv) && (iv != 0) != (hc.avClean != 0))
        throw std::runtime_error("restart: the file was written with avClean = " + std::to_string(iv) +
                                 ", the simulation was created with avClean = " + std::to_string(hc.avClean));
    RESTART_RC(sphb200_sim_set_constants(s, &hc));
    sphb200_sim_set_dt(s, info.minDt, info.minDt_m1);
    sphb200_sim_set_time(s, info.time);
    sphb200_sim_set_box(s, &info.box);

    sphb200_turb* t = sphb200_sim_turbulence(s);
    if (t && info.hasTurbulence)
    {
        int nm = sphb200_turb_num_modes(t);
        if (r->requireI64("turbulence::numModes") != nm)
            throw std::runtime_error("restart: mode table of the attached turbulence driver differs from the file");
        std::vector<double> phases(6 * size_t(nm));
        if (sphb200_restart_attribute(r, "turbulence::phases", phases.data(), phases.size()) != int64_t(phases.size()))
            throw std::runtime_error("restart: turbulence::phases has the wrong size");
        // TurbulenceData::loadOrStore (sph/hydro_turb/turbulence_data.hpp) reloads the forcing constants and the mode
        // table from the file; here the attached driver was built from the run's settings, so they must be the file's:
        // a driver with the same number of modes but another spectrum or Mach velocity would resume silently wrong
        {
            double have[4], want;
            std::vector<double> modes(3 * size_t(nm)), amps(1 * size_t(nm)), ph(6 * size_t(nm));
            std::vector<double> fileModes(modes.size()), fileAmps(amps.size());
            RESTART_RC(sphb200_turb_get(t, have, modes.data(), amps.data(), ph.data(), nullptr, nullptr));
            const char* names[4] = {"turbulence::variance", "turbulence::decayTime", "turbulence::solWeight",
                                    "turbulence::solWeightNorm"};
            for (int k = 0; k < 4; ++k)
                if (r->f64(names[k], &want) && want != have[k])
                    throw std::runtime_error(std::string("restart: ") + names[k] +
                                             " of the attached turbulence driver differs from the file");
            bool sameTable =
                sphb200_restart_attribute(r, "turbulence::modes", fileModes.data(), fileModes.size()) == int64_t(modes.size()) &&
                sphb200_restart_attribute(r, "turbulence::amplitudes", fileAmps.data(), fileAmps.size()) == int64_t(amps.size()) &&
                fileModes == modes && fileAmps == amps;
            if (!sameTable)
                throw std::runtime_error("restart: mode table of the attached turbulence driver differs from the file");
        }
        int64_t len = sphb200_restart_attribute(r, "rngEngineState", nullptr, 0);
        if (len <= 0) throw std::runtime_error("restart: rngEngineState missing");
        std::string rng(size_t(len), '\0');
        sphb200_restart_attribute(r, "rngEngineState", rng.data(), rng.size());
        RESTART_RC(sphb200_turb_set_state(t, phases.data(), rng.c_str()));
    }

    // HydroVeBdtProp::load (ve_hydro_bdt.hpp:155-168): optional, a dump of the global-time-step propagator has none
    const BlockHeader* dtm1 = r->find("ts::dt_m1", B_ATTRIBUTE);
    if (sphb200_sim_bdt_rungs(s) && dtm1 && r->i64("ts::numRungs", &iv))
    {
        if (dtm1->type != T_F32 || dtm1->count != 4) throw std::runtime_error("restart: ts::dt_m1 has the wrong shape");
        float dt_m1[4];
        r->read(dtm1->offset, dt_m1, sizeof(dt_m1));
        RESTART_RC(sphb200_sim_bdt_restore(s, int(iv), dt_m1));
    }
    RESTART_CATCH
}
