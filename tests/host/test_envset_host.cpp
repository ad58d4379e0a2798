// Drives the C++ host mirror (reinforcement-learning_b200/host/RLGymCPP/EnvSet.h) the way the reference's
// src/ExampleMain.cpp:66-140 builds its EnvSet: an EnvCreateFn returning arena + rewards + terminals + obs + parser + setter.
//   describe           CPU only: descriptor translation + loud failure on a user subclass (no device call)
//   run <meshdir> <ppt> <arenas> <steps>   GPU: ctor reset, StepFirstHalf(async) / Sync / StepSecondHalf, Reset; prints a summary line
#include "RLGymCPP/EnvSet.h"
#include <cmath>
#include <cstdlib>
#include <iostream>

using namespace RLGC;

static EnvCreateResult MakeEnv(int ppt, bool advanced, Reward* extra = nullptr) {
    EnvCreateResult r;
    r.arena = Arena::Create(GameMode::SOCCAR);
    for (int i = 0; i < ppt; i++) r.arena->AddCar(Team::BLUE);
    for (int i = 0; i < ppt; i++) r.arena->AddCar(Team::ORANGE);
    r.rewards = {{new VelocityPlayerToBallReward(), 1.f}, {new TouchBallReward(), 1}, {new GoalReward(), 10.f},
                 {new ZeroSumReward(new StrongTouchReward(20, 130), 0.5f), 2.f}};
    if (extra) r.rewards.push_back({extra, 1.f});
    r.terminalConditions = {new GoalScoreCondition(), new NoTouchCondition(10)};
    r.obsBuilder = advanced ? (ObsBuilder*)new AdvancedObs() : (ObsBuilder*)new DefaultObs();
    r.actionParser = new DefaultAction();
    r.stateSetter = new KickoffState();
    return r;
}

struct MyReward : Reward {};   // a user subclass the device knows nothing about

static int Describe() {
    EnvSetConfig cfg; cfg.numArenas = 4; cfg.tickSkip = 8; cfg.actionDelay = 7;
    ggl_envset_desc d = DescribeEnv(MakeEnv(2, true), cfg);
    bool ok = d.players_per_team == 2 && d.obs_kind == GGL_OBS_ADVANCED && d.num_rewards == 4 && d.num_terminals == 2 &&
              d.rewards[0].kind == GGL_REW_VELOCITY_PLAYER_TO_BALL && d.rewards[2].kind == GGL_REW_GOAL && d.rewards[2].params[0] == -1.f &&
              d.rewards[2].weight == 10.f && d.rewards[3].kind == GGL_REW_STRONG_TOUCH && d.rewards[3].zero_sum == 1 &&
              std::fabs(d.rewards[3].params[0] - 20 * 250.f / 9.f) < 1e-3f && d.terminals[1].kind == GGL_TERM_NO_TOUCH &&
              d.terminals[1].params[0] == 10.f && d.state_setter == GGL_SETTER_KICKOFF && d.tick_skip == 8 && d.action_delay == 7;
    {   // the random plugins of SURVEY 8 f4: DefaultObsPadded, CombinedState over { KickoffState, RandomState, FuzzedKickoffState }
        EnvCreateResult r = MakeEnv(2, false);
        r.obsBuilder = new DefaultObsPadded(3, Vec(1 / 5000.f, 1 / 6000.f, 1 / 2500.f), 1 / 3000.f, 0.25f);
        r.stateSetter = new CombinedState({{new KickoffState(), 1.f}, {new RandomState(true, false, true), 2.5f}, {new FuzzedKickoffState(), 0.5f}});
        ggl_envset_desc p = DescribeEnv(r, cfg);
        ok = ok && p.obs_kind == GGL_OBS_DEFAULT_PADDED && p.obs_max_players == 3 && p.state_setter == GGL_SETTER_COMBINED &&
             p.num_setter_children == 3 && p.setter_children[1].kind == GGL_SETTER_RANDOM &&
             p.setter_children[1].flags == (GGL_RANDOM_BALL_SPEED | GGL_RANDOM_CARS_ON_GROUND) && p.setter_children[1].weight == 2.5f &&
             p.setter_children[2].kind == GGL_SETTER_FUZZED_KICKOFF && p.obs_coefs[1] == 1 / 6000.f && p.obs_coefs[4] == 0.25f;
        ggl_envset_desc q = DescribeEnv(MakeEnv(2, false), cfg);        // stock DefaultObs(): its own default coefficients
        ok = ok && q.obs_coefs[0] == 1 / 4096.f && q.obs_coefs[3] == 1 / 2300.f && q.obs_coefs[4] == 1 / 5.5f;
        bool tooMany = false;
        r.obsBuilder = new DefaultObsPadded(1);
        try { DescribeEnv(r, cfg); } catch (const std::runtime_error&) { tooMany = true; }
        ok = ok && tooMany;
    }
    bool threw = false;
    try { DescribeEnv(MakeEnv(1, false, new MyReward()), cfg); } catch (const std::runtime_error& e) {
        threw = std::string(e.what()).find("no device implementation") != std::string::npos;
    }
    bool threwOrder = false;
    try {
        EnvCreateResult r = MakeEnv(1, false);
        r.arena->_cars.clear(); r.arena->AddCar(Team::ORANGE); r.arena->AddCar(Team::BLUE);
        DescribeEnv(r, cfg);
    } catch (const std::runtime_error&) { threwOrder = true; }
    std::cout << "describe ok=" << ok << " user_subclass_throws=" << threw << " roster_check=" << threwOrder << std::endl;
    return (ok && threw && threwOrder) ? 0 : 1;
}

static int Run(const char* meshDir, int ppt, int numArenas, int steps) {
    RocketSim::Init(meshDir);
    EnvSetConfig cfg;
    cfg.numArenas = numArenas; cfg.tickSkip = 8; cfg.actionDelay = 7; cfg.seed = 1;
    cfg.envCreateFn = [&](int) { return MakeEnv(ppt, true); };
    EnvSet env(cfg);
    const int np = env.state.numPlayers;
    if (np != numArenas * 2 * ppt || env.obsSize != 51 + 29 * 2 * ppt || env.numActions != 90) { std::cout << "bad sizes" << std::endl; return 1; }
    IList actions(np, 0);
    uint64_t rng = 12345;
    double rewSum = 0; int terminals = 0; bool finite = true;
    for (int s = 0; s < steps; s++) {
        env.Reset();
        env.StepFirstHalf(true);
        for (int i = 0; i < np; i++) {           // mask-valid random action from the host mirror of the masks
            const uint8_t* m = env.state.actionMasks.GetRow(i);
            int cnt = 0; for (int a = 0; a < 90; a++) cnt += m[a] != 0;
            rng = rng * 6364136223846793005ull + 1442695040888963407ull;
            int k = cnt ? (int)((rng >> 33) % cnt) : 0, sel = 0;
            for (int a = 0; a < 90; a++) if (m[a]) { if (k-- == 0) { sel = a; break; } }
            actions[i] = sel;
        }
        env.Sync();
        env.StepSecondHalf(actions, false);
        for (float v : env.state.obs.data) finite &= std::isfinite(v);
        for (float v : env.state.rewards) { finite &= std::isfinite(v); rewSum += v; }
        for (uint8_t t : env.state.terminals) terminals += t != 0;
    }
    ggl_arena_state st;
    env.GetArenaState(0, st);
    bool bad = false;
    try { IList wrong(np - 1, 0); env.StepSecondHalf(wrong, false); } catch (const std::runtime_error&) { bad = true; }
    std::cout << "run ok players=" << np << " obs=" << env.obsSize << " finite=" << finite << " reward_sum=" << rewSum
              << " terminals=" << terminals << " tick=" << st.tick_count << " ball_z=" << st.ball.pos[2] << " size_check=" << bad
              << " launches=" << ggl_launch_count() << std::endl;
    return (finite && bad && st.tick_count == (uint64_t)(8 * steps)) ? 0 : 1;
}

int main(int argc, char** argv) {
    if (argc >= 2 && std::string(argv[1]) == "describe") return Describe();
    if (argc >= 6 && std::string(argv[1]) == "run") return Run(argv[2], atoi(argv[3]), atoi(argv[4]), atoi(argv[5]));
    std::cerr << "usage: describe | run <meshdir> <ppt> <arenas> <steps>" << std::endl;
    return 2;
}
