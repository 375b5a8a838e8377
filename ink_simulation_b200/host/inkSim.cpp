// inkSim.cpp — the reference's run loop and .ply writers (src/inkSim.cpp:13-161) over the B200 `System`.
// Same frame pacing, same file names (<dir>/drop<i>/<frame>.ply, <dir>/velocityField.ply) and the same bytes:
// header text of src/inkSim.cpp:62-68 / :109-118 (including the non-standard "binary_small_endian" token, which
// render.py's reader does not look at) followed by raw host-endian fp32 records in particle index order.
// Differences, all on the I/O side: the drop<i>/ directories are created (the reference leaves that to the user and
// silently writes nothing otherwise, :49-51); a frame is ONE write() of a pinned buffer filled by one D2H copy instead
// of 3 ofstream::write calls per particle; the write overlaps the next solver steps; velocityField.ply holds exactly
// X*Y*Z records in i-j-k order (the reference dumps its hash map in bucket order, phantom cells included — SURVEY Q17).
#include "inkSim.h"

#include <chrono>
#include <cstdio>
#include <ctime>
#include <filesystem>
#include <stdexcept>

#include <cuda_runtime_api.h>

InkSim::InkSim(std::string writeDirectory) {
    this->ink_system.init();
    this->ink_system.checkNanAndInf();
    this->writeDirectory = writeDirectory;
}

InkSim::~InkSim() {
    drain();
    for (auto& b : m_buf)
        if (b.host) cudaFreeHost(b.host);
}

void InkSim::drain() {
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
    int currFrame = 1;
    float timeSinceLastRender = 0;
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
    const int drops = ink_system.numDrops();
    size_t total = 0;
    std::vector<size_t> off(drops + 1, 0);
    for (int i = 0; i < drops; i++) { total += (size_t)ink_system.numParticles(i) * 3; off[i + 1] = total; }
    FrameBuf& b = m_buf[m_cur];
    m_cur ^= 1;
    float* host = stage(b, total);
    std::vector<std::string> names(drops), headers(drops);
    for (int i = 0; i < drops; i++) {
        std::string path = this->writeDirectory + "/drop" + std::to_string(i);
        std::filesystem::create_directories(path);
        names[i] = path + "/" + std::to_string(frameNum) + ".ply";
        std::cout << "writing to: " + names[i] << std::endl;
        std::string numVertices = std::to_string(ink_system.numParticles(i));
        std::cout << "recording " + numVertices + " particles" << std::endl;
        headers[i] = particle_ply_header(ink_system.numParticles(i));
        ink_system.copyPositions(i, host + off[i]);   // D2H into pinned memory, original particle order
    }
    b.writer = std::thread([host, off, names, headers, drops]() {
        for (int i = 0; i < drops; i++) write_blob(names[i], headers[i], host + off[i], off[i + 1] - off[i]);
    });
}

void InkSim::writeWaterGridVelocities() {
    std::string filename = this->writeDirectory + "/" + "velocityField.ply";
    std::cout << "writing to: " + filename << std::endl;
    const size_t n = (size_t)WATERGRID_X * WATERGRID_Y * WATERGRID_Z;
    std::string numVectors = std::to_string(n);
    std::cout << "recording " + numVectors + " velocities" << std::endl;
    std::string header = "ply\n"
                         "format binary_small_endian 1.0\n"
                         "element vertex " + numVectors + "\n"
                         "property float x\n"
                         "property float y\n"
                         "property float z\n"
                         "property float u\n"
                         "property float v\n"
                         "property float w\n"
                         "end_header\n";
    std::vector<float> u(n), v(n), w(n), rec(6 * n);
    ink_system.copyVelocityField(u.data(), v.data(), w.data());
    const float offset = CELL_DIM / 2.f;
    size_t c = 0;
    for (int i = 0; i < WATERGRID_X; i++)
        for (int j = 0; j < WATERGRID_Y; j++)
            for (int k = 0; k < WATERGRID_Z; k++, c++) {
                float* r = &rec[6 * c];
                r[0] = i + offset; r[1] = j + offset; r[2] = k + offset;   // cell centre, src/inkSim.cpp:137-141
                r[3] = u[c]; r[4] = v[c]; r[5] = w[c];
            }
    write_blob(filename, header, rec.data(), rec.size());
}
