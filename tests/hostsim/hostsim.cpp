// TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT.
// Host (g++) compile of the device simulator headers under reinforcement-learning_b200/csrc so that the tick logic
// can be lock-stepped against the compiled reference (oracle/_ref) in a container WITHOUT a GPU.  The product
// library (libgigaenv_cuda.so) never links or loads this; it only launches CUDA kernels built from the same headers.
//   g++ -O2 -ffp-contract=off -shared -fPIC -I../../include hostsim.cpp -o _build/libhostsim.so
#include "../../reinforcement-learning_b200/csrc/sim_convert.cuh"
#include "../../reinforcement-learning_b200/csrc/env_logic.cuh"

using namespace ggl;

struct HostSim {
    WorldHost world;
    ArenaDev arena;
};

extern "C" {

void* hostsim_create(const ggl_mesh_blob* meshes, int n) {
    HostSim* h = new HostSim();
    for (int i = 0; i < n; i++)
        if (!world_add_cmf(h->world, meshes[i].data, meshes[i].size)) { delete h; return nullptr; }
    world_build_grid(h->world);
    memset(&h->arena, 0, sizeof(h->arena));
    h->arena.rng = 1;
    return h;
}
void hostsim_destroy(void* p) { delete (HostSim*)p; }
int hostsim_num_tris(void* p) { return (int)((HostSim*)p)->world.tris.size(); }

void hostsim_set_state(void* p, const ggl_arena_state* st) {
    HostSim* h = (HostSim*)p;
    arena_phys_from_host(*st, h->arena);
    for (int s = 0; s < h->arena.numCars; s++) h->arena.slotCar[s] = (uint8_t)(h->arena.numCars - 1 - s);
}
void hostsim_get_state(void* p, ggl_arena_state* st) { arena_phys_to_host(((HostSim*)p)->arena, *st); }
void hostsim_seed(void* p, uint32_t seed) { ((HostSim*)p)->arena.rng = seed; }

void hostsim_step(void* p, int ticks) {
    HostSim* h = (HostSim*)p;
    WorldDev W = world_view(h->world);
    ExecSerial ex;
    for (int i = 0; i < ticks; i++) arena_tick(h->arena, W, ex);
}

// contacts of the last tick: [n][20] like ref_arena_contacts (bodyA, bodyB, normal, posA, posB, dist, friction, restitution, special)
int hostsim_contacts(void* p, float* out, int max) {
    HostSim* h = (HostSim*)p;
    const ArenaDev& A = h->arena;
    int n = 0;
    for (int k = 0; k < A.numContacts && n < max; k++) {
        const Contact& c = A.contacts[k];
        float* o = out + 20 * n++;
        memset(o, 0, 80);
        o[0] = c.bodyA; o[1] = c.bodyB;
        o[2] = c.normal.x; o[3] = c.normal.y; o[4] = c.normal.z;
        o[11] = c.dist; o[12] = c.friction; o[13] = c.restitution; o[14] = c.special;
        o[5] = c.relA.x; o[6] = c.relA.y; o[7] = c.relA.z; o[8] = c.relB.x; o[9] = c.relB.y; o[10] = c.relB.z;
    }
    return n;
}


void hostsim_set_controls(void* p, const ggl_arena_state* st) {
    HostSim* h = (HostSim*)p;
    for (int i = 0; i < h->arena.numCars; i++) h->arena.cars[i].controls = controls_from_host(st->cars[i].controls);
}

size_t hostsim_sizeof_arena() { return sizeof(ArenaDev); }

}  // extern "C"
