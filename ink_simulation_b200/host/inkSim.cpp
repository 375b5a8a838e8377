// inkSim.cpp — the reference's run loop and .ply writers (src/inkSim.cpp:13-161) over the B200 `System`.
// Same frame pacing, same file names (<dir>/drop<i>/<frame>.ply, <dir>/velocityField.ply) and the same bytes:
// header text of src/inkSim.cpp:62-68 / :109-118 (including the non-standard "binary_small_endian" token, which
// render.py's reader does not look at) followed by raw host-endian fp32 records in particle index order.
// Differences, all on the I/O side: the drop<i>/ directories are created (the reference leaves that to the user and
// silently writes nothing otherwise, :49-51); a frame is ONE write() of a pinned buffer filled by one D2H copy instead
// of 3 ofstream::write calls per particle; both the copy (library copy stream, ink_fetch_positions_async) and the write (a
// writer thread) overlap the next solver steps; velocityField.ply holds exactly
// X*Y*Z records in i-j-k order (the reference dumps its hash map in bucket order, phantom cells included — SURVEY Q17).
#include "inkSim.h"

#include <chrono>
#include <cstdio>
#include <ctime>
#include <cstdint>
#include <cstring>
#include <filesystem>
#include <stdexcept>
#include <algorithm>

#include <cuda_runtime_api.h>

static void write_blob(const std::string& filename, const std::string& header, const float* data, size_t floats);

InkSim::InkSim(std::string writeDirectory) {
    this->ink_system.init();
    this->ink_system.checkNanAndInf();
    this->writeDirectory = writeDirectory;
}

InkSim::~InkSim() {
    try { drain(); } catch (const std::exception& e) { std::cerr << "InkSim: " << e.what() << std::endl; }   // never throw from a destructor
    for (auto& b : m_buf)
        if (b.host) cudaFreeHost(b.host);
}

void InkSim::flushPending() {
    bool any = false;
    for (auto& b : m_buf) any = any || b.pending;
    if (!any) return;
    ink_system.syncIO();   // every enqueued device-to-host copy has landed
    for (auto& b : m_buf) {
        if (!b.pending) continue;
        b.pending = false;
        float* host = b.host;
        b.writer = std::thread([host, off = b.off, names = b.names, headers = b.headers]() {
            for (size_t i = 0; i + 1 < off.size(); i++) write_blob(names[i], headers[i], host + off[i], off[i + 1] - off[i]);
        });
    }
}

void InkSim::drain() {
    flushPending();
    for (auto& b : m_buf)
        if (b.writer.joinable()) b.writer.join();
}

float* InkSim::stage(FrameBuf& b, size_t floats) {
    if (b.writer.joinable()) b.writer.join();   // the buffer is free once its previous frame is on disk
    if (floats > b.cap) {
        if (b.host) cudaFreeHost(b.host);
        if (cudaMallocHost((void**)&b.host, floats * sizeof(float)) != cudaSuccess) throw std::runtime_error("cudaMallocHost failed");
        b.cap = floats;
    }
    return b.host;
}

void InkSim::simulate(const float renderTimestep, int totalTimesteps) {
    auto startTS = std::chrono::system_clock::now();
    auto startTime = std::chrono::system_clock::to_time_t(startTS);
    std::cout << "\033[32mSimulation started at: \033[m" << std::ctime(&startTime) << std::endl;

    int timestepCounter = 1;
    int currFrame = m_resumed ? m_nextFrame : 1;
    float timeSinceLastRender = m_resumed ? m_timeSinceLastRender : 0;
    m_resumed = false;
    while (timestepCounter <= totalTimesteps) {
        float timeStep = this->ink_system.solve(renderTimestep - timeSinceLastRender);
        timeSinceLastRender += timeStep;
        if (timeSinceLastRender >= renderTimestep) {
            writeToFile(currFrame);
            currFrame += 1;
            timeSinceLastRender = 0;
        }
        timestepCounter += 1;
    }
    drain();
    writeWaterGridVelocities();
    m_nextFrame = currFrame;
    m_timeSinceLastRender = timeSinceLastRender;

    auto endTS = std::chrono::system_clock::now();
    std::chrono::duration<double> elapsed_seconds = endTS - startTS;
    int seconds = elapsed_seconds.count();
    std::cout << "\033[32mSimulation & Writing took " << seconds / 60 << " min " << seconds % 60 << " sec\033[m" << std::endl;
}

static void write_blob(const std::string& filename, const std::string& header, const float* data, size_t floats) {
    FILE* f = std::fopen(filename.c_str(), "wb");
    if (!f) { std::cerr << "cannot open " << filename << std::endl; return; }
    std::fwrite(header.data(), 1, header.size(), f);
    std::fwrite(data, sizeof(float), floats, f);
    std::fclose(f);
}

static std::string particle_ply_header(long long n) {
    return "ply\n"
           "format binary_small_endian 1.0\n"
           "element vertex " + std::to_string(n) + "\n"
           "property float x\n"
           "property float y\n"
           "property float z\n"
           "end_header\n";
}

// C hook for the CPU-side tests: the frame writer on caller-provided positions (no GPU context involved)
extern "C" void inkhost_write_particles_ply(const char* filename, const float* xyz, long n) {
    write_blob(filename, particle_ply_header(n), xyz, (size_t)n * 3);
}

void InkSim::writeToFile(int frameNum) {
    flushPending();   // the previous frame's copy has had the steps in between to complete; its file write starts now
    const int drops = ink_system.numDrops();
    size_t total = 0;
    std::vector<size_t> off(drops + 1, 0);
    for (int i = 0; i < drops; i++) { total += (size_t)ink_system.numParticles(i) * 3; off[i + 1] = total; }
    FrameBuf& b = m_buf[m_cur];
    m_cur ^= 1;
    float* host = stage(b, total);   // waits until the frame before the previous one is on disk: its buffer is reused
    b.off = off;
    b.names.assign(drops, "");
    b.headers.assign(drops, "");
    for (int i = 0; i < drops; i++) {
        std::string path = this->writeDirectory + "/drop" + std::to_string(i);
        std::filesystem::create_directories(path);
        b.names[i] = path + "/" + std::to_string(frameNum) + ".ply";
        std::cout << "writing to: " + b.names[i] << std::endl;
        std::string numVertices = std::to_string(ink_system.numParticles(i));
        std::cout << "recording " + numVertices + " particles" << std::endl;
        b.headers[i] = particle_ply_header(ink_system.numParticles(i));
        ink_system.fetchPositionsAsync(i, host + off[i]);   // snapshot in stream order, D2H on the copy stream, original particle order
    }
    b.pending = true;
}

// header of a dump with `props` float properties per record; the first three are always x y z
static std::string ply_header(size_t n, const std::vector<std::string>& props) {
    std::string h = "ply\n"
                    "format binary_small_endian 1.0\n"
                    "element vertex " + std::to_string(n) + "\n";
    for (const std::string& p : props) h += "property float " + p + "\n";
    return h + "end_header\n";
}

// (cx, cy, cz, f0[, f1, f2]) per cell in i-j-k order, centre = index + h/2 (src/inkSim.cpp:137-141)
static void write_grid_ply(const std::string& filename, const std::vector<std::string>& names, const std::vector<const float*>& f) {
    const size_t n = (size_t)WATERGRID_X * WATERGRID_Y * WATERGRID_Z, props = 3 + f.size();
    std::vector<std::string> all = {"x", "y", "z"};
    all.insert(all.end(), names.begin(), names.end());
    std::vector<float> rec(props * n);
    const float offset = CELL_DIM / 2.f;
    size_t c = 0;
    for (int i = 0; i < WATERGRID_X; i++)
        for (int j = 0; j < WATERGRID_Y; j++)
            for (int k = 0; k < WATERGRID_Z; k++, c++) {
                float* r = &rec[props * c];
                r[0] = i + offset; r[1] = j + offset; r[2] = k + offset;
                for (size_t a = 0; a < f.size(); ++a) r[3 + a] = f[a][c];
            }
    write_blob(filename, ply_header(n, all), rec.data(), rec.size());
}

// Reader for every .ply this program writes (frames, velocityField.ply, checkpoint files): like render.py:290-294 it only
// looks for "element vertex N" and "end_header", plus the number of "property float" lines; the payload is raw host-endian
// fp32.  Returns the records (n x props) or throws.
static std::vector<float> read_ply(const std::string& filename, size_t& n, size_t& props) {
    FILE* f = std::fopen(filename.c_str(), "rb");
    if (!f) throw std::runtime_error("cannot open " + filename);
    std::string header;
    n = 0; props = 0;
    bool ended = false, have_n = false;
    char line[512];
    while (std::fgets(line, sizeof line, f)) {
        const std::string ln(line);
        if (ln.rfind("element vertex ", 0) == 0) { n = (size_t)std::stoull(ln.substr(15)); have_n = true; }
        else if (ln.rfind("property float ", 0) == 0) props++;
        else if (ln == "end_header\n") { ended = true; break; }
    }
    if (!ended || !have_n || props == 0) { std::fclose(f); throw std::runtime_error("not a ply dump of this program: " + filename); }
    std::vector<float> rec(n * props);
    const size_t got = std::fread(rec.data(), sizeof(float), rec.size(), f);
    const bool extra = std::fgetc(f) != EOF;
    std::fclose(f);
    if (got != rec.size() || extra) throw std::runtime_error("payload size does not match the header: " + filename);
    return rec;
}

// C hooks for the CPU-side tests (no GPU context involved)
extern "C" long inkhost_read_ply(const char* filename, float* out, long cap_floats, int* props_out) {
    try {
        size_t n, props;
        const std::vector<float> rec = read_ply(filename, n, props);
        if (props_out) *props_out = (int)props;
        if (out && (long)rec.size() <= cap_floats) std::copy(rec.begin(), rec.end(), out);
        return (long)n;
    } catch (const std::exception&) { return -1; }
}
extern "C" void inkhost_write_grid_ply(const char* filename, int nfields, const float* f0, const float* f1, const float* f2) {
    static const char* vel[3] = {"u", "v", "w"};
    std::vector<std::string> names;
    std::vector<const float*> f;
    const float* in[3] = {f0, f1, f2};
    for (int a = 0; a < nfields && a < 3; ++a) { names.push_back(nfields == 1 ? "q" : vel[a]); f.push_back(in[a]); }
    write_grid_ply(filename, names, f);
}
extern "C" void inkhost_grid_dims(int* xyz) { xyz[0] = WATERGRID_X; xyz[1] = WATERGRID_Y; xyz[2] = WATERGRID_Z; }

void InkSim::writeWaterGridVelocities() {
    std::string filename = this->writeDirectory + "/" + "velocityField.ply";
    std::cout << "writing to: " + filename << std::endl;
    const size_t n = (size_t)WATERGRID_X * WATERGRID_Y * WATERGRID_Z;
    std::cout << "recording " + std::to_string(n) + " velocities" << std::endl;
    std::vector<float> u(n), v(n), w(n);
    ink_system.copyVelocityField(u.data(), v.data(), w.data());
    write_grid_ply(filename, {"u", "v", "w"}, {u.data(), v.data(), w.data()});
}

// ---- checkpoint / resume ------------------------------------------------------------------------------------------------
// State of a run between two steps = currVelocity (old == curr there, src/waterSolver.cpp:92-95), the particles, and — ours
// only — the start vector of the next pressure solve and the phantom-cell velocity (Q8; meta.txt, as double bits).
// Payloads are the reference's fp32 .ply: a precision = 64 context resumes to fp32 rounding, not bit for bit.  Everything else the library keeps (force-mask seeds, max speed,
// particle bins, curl) is recomputed from these, and none of it depends on the particle order in memory.
void InkSim::writeCheckpoint() {
    drain();
    const std::string dir = this->writeDirectory + "/checkpoint";
    std::filesystem::create_directories(dir);
    const size_t n = (size_t)WATERGRID_X * WATERGRID_Y * WATERGRID_Z;
    std::vector<float> u(n), v(n), w(n), q(n);
    ink_system.copyVelocityField(u.data(), v.data(), w.data());
    write_grid_ply(dir + "/velocityField.ply", {"u", "v", "w"}, {u.data(), v.data(), w.data()});
    ink_system.copyWarmStart(q.data());
    write_grid_ply(dir + "/warmstart.ply", {"q"}, {q.data()});
    const int drops = ink_system.numDrops();
    for (int d = 0; d < drops; d++) {
        const size_t np = (size_t)ink_system.numParticles(d);
        std::vector<float> pos(3 * np), vel(3 * np), rec(6 * np);
        if (np) ink_system.copyParticles(d, pos.data(), vel.data());
        for (size_t i = 0; i < np; i++)
            for (int a = 0; a < 3; a++) { rec[6 * i + a] = pos[3 * i + a]; rec[6 * i + 3 + a] = vel[3 * i + a]; }
        write_blob(dir + "/drop" + std::to_string(d) + ".ply", ply_header(np, {"x", "y", "z", "vx", "vy", "vz"}), rec.data(), rec.size());
    }
    uint32_t bits;
    static_assert(sizeof bits == sizeof m_timeSinceLastRender, "float bits");
    std::memcpy(&bits, &m_timeSinceLastRender, sizeof bits);
    FILE* f = std::fopen((dir + "/meta.txt").c_str(), "w");
    if (!f) throw std::runtime_error("cannot write " + dir + "/meta.txt");
    std::fprintf(f, "grid %d %d %d\ndrops %d\nnext_frame %d\ntime_since_last_render_bits 0x%08x\n", WATERGRID_X, WATERGRID_Y, WATERGRID_Z, drops,
                 m_nextFrame, bits);
    // phantom cells of out-of-grid particles: their accumulated velocity enters calcTimeStep's maximum
    double pv[3] = {0, 0, 0};
    const int have = ink_system.getPhantom(pv) ? 1 : 0;
    unsigned long long pb[3];
    std::memcpy(pb, pv, sizeof pb);
    std::fprintf(f, "phantom %d 0x%016llx 0x%016llx 0x%016llx\n", (int)have, pb[0], pb[1], pb[2]);
    std::fclose(f);
}

bool InkSim::resume() {
    const std::string dir = this->writeDirectory + "/checkpoint";
    FILE* f = std::fopen((dir + "/meta.txt").c_str(), "r");
    if (!f) return false;
    int X = 0, Y = 0, Z = 0, drops = 0, frame = 1;
    unsigned int bits = 0;
    const int got = std::fscanf(f, "grid %d %d %d drops %d next_frame %d time_since_last_render_bits 0x%x", &X, &Y, &Z, &drops, &frame, &bits);
    int have = 0;
    unsigned long long pb[3] = {0, 0, 0};
    const bool got_phantom = (got == 6) && std::fscanf(f, " phantom %d 0x%llx 0x%llx 0x%llx", &have, &pb[0], &pb[1], &pb[2]) == 4;   // absent in older checkpoints
    std::fclose(f);
    if (got != 6) throw std::runtime_error("malformed " + dir + "/meta.txt");
    if (X != WATERGRID_X || Y != WATERGRID_Y || Z != WATERGRID_Z) throw std::runtime_error("checkpoint is for another grid size");
    const size_t n = (size_t)X * Y * Z;
    size_t cnt, props;
    std::vector<float> rec = read_ply(dir + "/velocityField.ply", cnt, props);
    if (cnt != n || props != 6) throw std::runtime_error("checkpoint velocityField.ply: wrong shape");
    std::vector<float> u(n), v(n), w(n), q(n);
    for (size_t c = 0; c < n; c++) { u[c] = rec[6 * c + 3]; v[c] = rec[6 * c + 4]; w[c] = rec[6 * c + 5]; }
    rec = read_ply(dir + "/warmstart.ply", cnt, props);
    if (cnt != n || props != 4) throw std::runtime_error("checkpoint warmstart.ply: wrong shape");
    for (size_t c = 0; c < n; c++) q[c] = rec[4 * c + 3];
    std::vector<std::vector<float>> pos(drops), vel(drops);
    for (int d = 0; d < drops; d++) {   // read everything before touching the state
        rec = read_ply(dir + "/drop" + std::to_string(d) + ".ply", cnt, props);
        if (props != 6) throw std::runtime_error("checkpoint drop file: wrong shape");
        pos[d].resize(3 * cnt); vel[d].resize(3 * cnt);
        for (size_t i = 0; i < cnt; i++)
            for (int a = 0; a < 3; a++) { pos[d][3 * i + a] = rec[6 * i + a]; vel[d][3 * i + a] = rec[6 * i + 3 + a]; }
    }
    ink_system.setVelocityField(u.data(), v.data(), w.data());
    ink_system.setWarmStart(q.data());
    for (int d = 0; d < drops; d++) ink_system.setInk(d, pos[d].data(), vel[d].data(), (long long)(pos[d].size() / 3));
    if (got_phantom) {   // after setInk: seeding the force mask from the restored particles must not redefine it
        double pv[3];
        std::memcpy(pv, pb, sizeof pv);
        ink_system.setPhantom(have != 0, pv);
    }
    float since;
    std::memcpy(&since, &bits, sizeof since);
    m_nextFrame = frame;
    m_timeSinceLastRender = since;
    m_resumed = true;
    return true;
}
