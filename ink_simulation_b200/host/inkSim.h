// inkSim.h — drop-in for the reference's src/inkSim.h: same class, same constructor, same simulate() signature.
#ifndef INK_SIM_H
#define INK_SIM_H
#include <string>
#include <thread>
#include <vector>

#include "system.h"

class InkSim {
public:
    /// @param writeDirectory the path to the *directory* to save ply files to
    InkSim(std::string writeDirectory);
    ~InkSim();

    /// run the ink simulation: a frame is written every `renderTimestep` of simulated time, `totalTimesteps` solver steps
    void simulate(const float renderTimestep, int totalTimesteps);

    // ---- additions (no reference equivalent: the reference can dump velocityField.ply but never read it back) ------------
    /// write <dir>/checkpoint/: velocityField.ply (the reference's dump format), warmstart.ply (x y z q), drop<i>.ply
    /// (x y z vx vy vz, particle index order) and meta.txt (next frame number, time since the last frame as float bits)
    void writeCheckpoint();
    /// restore what writeCheckpoint() wrote; the next simulate() continues the frame numbering and pacing from there,
    /// bit for bit as if the run had never stopped.  Returns false (state untouched) when there is no checkpoint.
    bool resume();

private:
    System ink_system;
    std::string writeDirectory;

    void writeToFile(int frameNum);
    void writeWaterGridVelocities();

    // asynchronous frame output: positions land in one of two pinned host buffers, a writer thread streams the
    // previous frame to disk while the GPU already computes the next steps
    // — and the device-to-host copy of a frame runs on the library's copy stream while the next steps compute: a frame's
    // copy is only waited for (and its file write started) when the NEXT frame is due, at a checkpoint or at the end
    struct FrameBuf {
        float* host = nullptr; size_t cap = 0; std::thread writer;
        bool pending = false;                       // copy enqueued, file write not started yet
        std::vector<size_t> off; std::vector<std::string> names, headers;
    };
    // frame pacing carried across simulate() calls only through resume(): like the reference, a plain second call to
    // simulate() starts again at frame 1
    int m_nextFrame = 1;
    float m_timeSinceLastRender = 0;
    bool m_resumed = false;

    FrameBuf m_buf[2];
    int m_cur = 0;
    float* stage(FrameBuf& b, size_t floats);
    void flushPending();                            // wait for enqueued copies, hand their frames to writer threads
    void drain();                                   // flushPending + wait for every file to be on disk
};

#endif  // INK_SIM_H
