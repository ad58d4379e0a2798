// Host-side mirror of the RLGC plugin surface for the device-resident EnvSet (C++17, header-only, no torch, no CUDA
// headers: everything goes through the C ABI in include/gigaenv.h -> libgigaenv_cuda.so).
//
// What it mirrors (paths relative to /root/reference/GigaLearnCPP/RLGymCPP/src/RLGymCPP unless noted):
//   RLGC::EnvCreateResult / EnvCreateFn / EnvSetConfig / EnvState / EnvSet      EnvSet/EnvSet.h:14-110
//   RLGC::Reward / WeightedReward / ZeroSumReward                               Rewards/Reward.h:7-90, ZeroSumReward.h:7-33
//   the stock rewards                                                           Rewards/CommonRewards.h:7-203
//   RLGC::TerminalCondition, GoalScoreCondition, NoTouchCondition               TerminalConditions/*.h
//   RLGC::ObsBuilder, DefaultObs, AdvancedObs                                   ObsBuilders/*.h
//   RLGC::ActionParser, DefaultAction                                           ActionParsers/DefaultAction.h
//   RLGC::StateSetter, KickoffState                                             StateSetters/KickoffState.h
//   RocketSim::Init / InitFromMem, Arena::Create, Arena::AddCar, Team, GameMode RocketSim/src/RocketSim.cpp:70-175, Sim/Arena/Arena.h:111
//
// The reference runs plugins as virtual calls per player per step on the CPU. Here a plugin object is a DESCRIPTOR: every
// stock class knows how to describe itself to the device (Describe()), and the EnvSet constructor translates the
// EnvCreateResult of arena 0 into one ggl_envset_desc (all arenas must be configured identically, which it checks by
// describing every arena's plugins). A user subclass that the device has no implementation for makes the constructor
// throw std::runtime_error exactly like RG_ERR_CLOSE (Framework.h:16-21) -- there is NO CPU fallback.
//
// Threading contract (EnvSet.cpp:113-130, Learner.cpp:701-793): StepFirstHalf(true) returns at once (the launch is
// asynchronous on the envset's stream), Sync() waits, StepSecondHalf(actions, false) returns with `state` filled.
// state.obs / actionMasks / rewards / terminals are HOST mirrors filled after StepSecondHalf and Reset; the device
// buffers libtorch would wrap zero-copy are in deviceViews().
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <functional>
#include <stdexcept>
#include <string>
#include <vector>
#include <dirent.h>

#include "../../../include/gigaenv.h"

// ---------------------------------------------------------------------------------------------- RocketSim stand-ins
enum class GameMode : uint8_t { SOCCAR = 0, HOOPS, HEATSEEKER, SNOWDAY, THE_VOID };   // Sim/GameMode.h:6-16; the device integrates SOCCAR and HEATSEEKER
enum class Team : uint8_t { BLUE = 0, ORANGE = 1 };
struct CarConfig {};
static const CarConfig CAR_CONFIG_OCTANE = {};

namespace RocketSim {
// Collision meshes: the reference loads every *.cmf under <folder>/soccar (RocketSim.cpp:70-100) or takes them from
// memory (InitFromMem, :102-175). The blobs are kept here and handed to every EnvSet created afterwards.
inline std::vector<std::vector<uint8_t>>& _Meshes() { static std::vector<std::vector<uint8_t>> m; return m; }
inline void InitFromMem(const std::vector<std::vector<uint8_t>>& meshFiles) { _Meshes() = meshFiles; }
inline void Init(const std::string& collisionMeshesFolder) {
    std::vector<std::vector<uint8_t>> files;
    const std::string dir = collisionMeshesFolder + "/soccar";
    DIR* d = opendir(dir.c_str());
    if (!d) throw std::runtime_error("RocketSim::Init: cannot open " + dir);
    std::vector<std::string> names;
    while (dirent* e = readdir(d)) {
        std::string n = e->d_name;
        if (n.size() > 4 && n.substr(n.size() - 4) == ".cmf") names.push_back(n);
    }
    closedir(d);
    std::sort(names.begin(), names.end());
    for (auto& n : names) {
        std::ifstream f(dir + "/" + n, std::ios::binary);
        files.emplace_back((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    }
    if (files.empty()) throw std::runtime_error("RocketSim::Init: no .cmf files in " + dir);
    _Meshes() = std::move(files);
}
}  // namespace RocketSim

struct Car { uint32_t id; Team team; };
struct Vec {                                            // RocketSim Math/MathTypes/MathTypes.h:8-17 (what plugin constructors take)
    float x = 0, y = 0, z = 0;
    Vec() = default;
    Vec(float x, float y, float z) : x(x), y(y), z(z) {}
};

// Arena::Create + AddCar record the roster; the physics state lives on the device (one blob per arena).
struct Arena {
    GameMode gameMode;
    std::vector<Car> _cars;
    static Arena* Create(GameMode mode) { return new Arena{mode, {}}; }
    Car* AddCar(Team team, const CarConfig& = CAR_CONFIG_OCTANE) {
        _cars.push_back(Car{(uint32_t)_cars.size() + 1, team});
        return &_cars.back();
    }
};

namespace RLGC {

typedef std::vector<float> FList;
typedef std::vector<int> IList;

[[noreturn]] inline void _Fatal(const std::string& what) { throw std::runtime_error("RLGC(gigaenv-b200) FATAL ERROR: " + what); }

template <typename T>
struct DimList2 {                                       // BasicTypes/Lists.h:33-75
    size_t size[2] = {0, 0};
    size_t numel = 0;
    std::vector<T> data;
    void Resize(size_t a, size_t b) { size[0] = a; size[1] = b; numel = a * b; data.resize(numel); }
    T& At(size_t i, size_t j) { return data[i * size[1] + j]; }
    const T* GetRow(size_t i) const { return data.data() + i * size[1]; }
};

// ---------------------------------------------------------------------------------------------- plugins = descriptors
class Reward {                                          // Rewards/Reward.h:7-82
public:
    // fills kind/params; false = "no device implementation" (user subclass)
    virtual bool Describe(ggl_reward_desc&) const { return false; }
    virtual std::string GetName() const { return "Reward"; }
    virtual ~Reward() {}
};
struct WeightedReward {                                 // Rewards/Reward.h:84-90
    Reward* reward; float weight;
    WeightedReward(Reward* r, float w) : reward(r), weight(w) {}
    WeightedReward(Reward* r, int w) : reward(r), weight((float)w) {}
};

#define GGL_STOCK_REWARD(Name, Kind)                                                                   \
    class Name : public Reward { public:                                                               \
        bool Describe(ggl_reward_desc& d) const override { d.kind = Kind; return true; }               \
        std::string GetName() const override { return #Name; } };
GGL_STOCK_REWARD(PlayerGoalReward, GGL_REW_PLAYER_GOAL)          // CommonRewards.h:20-28
GGL_STOCK_REWARD(AssistReward, GGL_REW_ASSIST)
GGL_STOCK_REWARD(ShotReward, GGL_REW_SHOT)
GGL_STOCK_REWARD(ShotPassReward, GGL_REW_SHOT_PASS)
GGL_STOCK_REWARD(SaveReward, GGL_REW_SAVE)
GGL_STOCK_REWARD(BumpReward, GGL_REW_BUMP)
GGL_STOCK_REWARD(BumpedPenalty, GGL_REW_BUMPED_PENALTY)
GGL_STOCK_REWARD(DemoReward, GGL_REW_DEMO)
GGL_STOCK_REWARD(DemoedPenalty, GGL_REW_DEMOED_PENALTY)
GGL_STOCK_REWARD(VelocityPlayerToBallReward, GGL_REW_VELOCITY_PLAYER_TO_BALL)   // :75-82
GGL_STOCK_REWARD(FaceBallReward, GGL_REW_FACE_BALL)                             // :85-91
GGL_STOCK_REWARD(TouchBallReward, GGL_REW_TOUCH_BALL)                           // :93-98
GGL_STOCK_REWARD(SpeedReward, GGL_REW_SPEED)                                    // :100-105
GGL_STOCK_REWARD(WavedashReward, GGL_REW_WAVEDASH)                              // :107-119
GGL_STOCK_REWARD(PickupBoostReward, GGL_REW_PICKUP_BOOST)                       // :121-133
GGL_STOCK_REWARD(AirReward, GGL_REW_AIR)                                        // :147-152
GGL_STOCK_REWARD(TouchAccelReward, GGL_REW_TOUCH_ACCEL)                         // :157-178
#undef GGL_STOCK_REWARD

class GoalReward : public Reward { public:              // :32-44
    float concedeScale;
    GoalReward(float concedeScale = -1) : concedeScale(concedeScale) {}
    bool Describe(ggl_reward_desc& d) const override { d.kind = GGL_REW_GOAL; d.params[0] = concedeScale; return true; }
    std::string GetName() const override { return "GoalReward"; } };
class VelocityReward : public Reward { public:          // :47-54
    bool isNegative;
    VelocityReward(bool isNegative = false) : isNegative(isNegative) {}
    bool Describe(ggl_reward_desc& d) const override { d.kind = GGL_REW_VELOCITY; d.params[0] = isNegative ? 1.f : 0.f; return true; }
    std::string GetName() const override { return "VelocityReward"; } };
class VelocityBallToGoalReward : public Reward { public: // :57-72
    bool ownGoal;
    VelocityBallToGoalReward(bool ownGoal = false) : ownGoal(ownGoal) {}
    bool Describe(ggl_reward_desc& d) const override { d.kind = GGL_REW_VELOCITY_BALL_TO_GOAL; d.params[0] = ownGoal ? 1.f : 0.f; return true; }
    std::string GetName() const override { return "VelocityBallToGoalReward"; } };
class SaveBoostReward : public Reward { public:         // :136-144
    float exponent;
    SaveBoostReward(float exponent = 0.5f) : exponent(exponent) {}
    bool Describe(ggl_reward_desc& d) const override { d.kind = GGL_REW_SAVE_BOOST; d.params[0] = exponent; return true; }
    std::string GetName() const override { return "SaveBoostReward"; } };
class StrongTouchReward : public Reward { public:       // :180-202 (KPH -> uu/s: Math.h:12-14)
    float minRewardedVel, maxRewardedVel;
    StrongTouchReward(float minSpeedKPH = 20, float maxSpeedKPH = 130)
        : minRewardedVel(minSpeedKPH * (250.f / 9.f)), maxRewardedVel(maxSpeedKPH * (250.f / 9.f)) {}
    bool Describe(ggl_reward_desc& d) const override {
        d.kind = GGL_REW_STRONG_TOUCH; d.params[0] = minRewardedVel; d.params[1] = maxRewardedVel; return true; }
    std::string GetName() const override { return "StrongTouchReward"; } };

class KickoffProximityReward2v2Enhanced : public Reward { public:     // Rewards/KickoffProximityReward2v2Enhanced.h:7-367
    float goerReward = 1.2f, cheaterReward = 0.6f, dynamicWeight = 0.3f, rotationPrepWeight = 0.2f;
    bool Describe(ggl_reward_desc& d) const override {
        d.kind = GGL_REW_KICKOFF_PROXIMITY_2V2;
        d.params[0] = goerReward; d.params[1] = cheaterReward; d.params[2] = dynamicWeight; d.params[3] = rotationPrepWeight;
        return true; }
    std::string GetName() const override { return "KickoffProximityReward2v2Enhanced"; } };

// ZeroSumReward (ZeroSumReward.h:7-33). In the EnvSet path the reference calls GetReward() per player, which the wrapper
// forwards to its child untouched (confirmed by running the compiled reference): a pass-through.
class ZeroSumReward : public Reward { public:
    Reward* child; float teamSpirit, opponentScale;
    ZeroSumReward(Reward* child, float teamSpirit, float opponentScale = 1, bool = true)
        : child(child), teamSpirit(teamSpirit), opponentScale(opponentScale) {}
    ~ZeroSumReward() override { delete child; }         // RewardWrapper owns its child (RewardWrapper.h:11-17)
    bool Describe(ggl_reward_desc& d) const override { if (!child || !child->Describe(d)) return false; d.zero_sum = 1; return true; }
    std::string GetName() const override { return child ? child->GetName() : "ZeroSumReward"; } };

class TerminalCondition {                               // TerminalConditions/TerminalCondition.h:13-21
public:
    virtual bool Describe(ggl_terminal_desc&) const { return false; }
    virtual bool IsTruncation() { return false; }
    virtual ~TerminalCondition() {}
};
class GoalScoreCondition : public TerminalCondition { public:   // GoalScoreCondition.h:8-14
    bool Describe(ggl_terminal_desc& d) const override { d.kind = GGL_TERM_GOAL_SCORE; return true; } };
class NoTouchCondition : public TerminalCondition { public:     // NoTouchCondition.h:5-33
    float maxTime;
    NoTouchCondition(float maxTime) : maxTime(maxTime) {}
    bool Describe(ggl_terminal_desc& d) const override { d.kind = GGL_TERM_NO_TOUCH; d.params[0] = maxTime; return true; }
    bool IsTruncation() override { return true; } };

class ObsBuilder { public:                              // ObsBuilders/ObsBuilder.h:8-14
    virtual int DescribeKind() const { return -1; }
    virtual int DescribeMaxPlayers() const { return 0; }
    virtual void DescribeCoefs(float out[5]) const { for (int i = 0; i < 5; i++) out[i] = 0.f; }    // all 0 = the stock defaults
    virtual ~ObsBuilder() {} };
namespace CommonValues {                                // CommonValues.h:20-40 (the values the obs defaults are built from)
constexpr float SIDE_WALL_X = 4096, BACK_WALL_Y = 5120, CEILING_Z = 2044, CAR_MAX_SPEED = 2300, CAR_MAX_ANG_VEL = 5.5f;
}
class DefaultObs : public ObsBuilder { public:          // ObsBuilders/DefaultObs.h:6-22
    Vec posCoef; float velCoef, angVelCoef;
    DefaultObs(Vec posCoef = Vec(1 / CommonValues::SIDE_WALL_X, 1 / CommonValues::BACK_WALL_Y, 1 / CommonValues::CEILING_Z),
               float velCoef = 1 / CommonValues::CAR_MAX_SPEED, float angVelCoef = 1 / CommonValues::CAR_MAX_ANG_VEL)
        : posCoef(posCoef), velCoef(velCoef), angVelCoef(angVelCoef) {}
    int DescribeKind() const override { return GGL_OBS_DEFAULT; }
    void DescribeCoefs(float out[5]) const override { out[0] = posCoef.x; out[1] = posCoef.y; out[2] = posCoef.z; out[3] = velCoef; out[4] = angVelCoef; } };
class AdvancedObs : public ObsBuilder { public: int DescribeKind() const override { return GGL_OBS_ADVANCED; } };
class DefaultObsPadded : public DefaultObs { public:    // ObsBuilders/DefaultObsPadded.h:9-24
    int maxPlayers;
    DefaultObsPadded(int maxPlayers, Vec posCoef = Vec(1 / CommonValues::SIDE_WALL_X, 1 / CommonValues::BACK_WALL_Y, 1 / CommonValues::CEILING_Z),
                     float velCoef = 1 / CommonValues::CAR_MAX_SPEED, float angVelCoef = 1 / CommonValues::CAR_MAX_ANG_VEL)
        : DefaultObs(posCoef, velCoef, angVelCoef), maxPlayers(maxPlayers) {}
    int DescribeKind() const override { return GGL_OBS_DEFAULT_PADDED; }
    int DescribeMaxPlayers() const override { return maxPlayers; } };

class ActionParser { public:                            // ActionParsers/ActionParser.h:9-19
    virtual bool IsDefaultAction() const { return false; }
    virtual int GetActionAmount() const { return 0; }
    virtual ~ActionParser() {} };
class DefaultAction : public ActionParser { public:     // DefaultAction.h:8-24 (90 actions)
    bool IsDefaultAction() const override { return true; }
    int GetActionAmount() const override { return GGL_NUM_ACTIONS; } };

class StateSetter { public:                             // StateSetters/StateSetter.h:5-8
    virtual int DescribeKind() const { return -1; }
    virtual int DescribeFlags() const { return 0; }
    virtual ~StateSetter() {} };
class KickoffState : public StateSetter { public: int DescribeKind() const override { return GGL_SETTER_KICKOFF; } };
// Reset states supplied by the host (EnvSet::SetResetState): the route for user StateSetters whose output is data.
class InjectedState : public StateSetter { public: int DescribeKind() const override { return GGL_SETTER_INJECTED; } };
class FuzzedKickoffState : public StateSetter { public: int DescribeKind() const override { return GGL_SETTER_FUZZED_KICKOFF; } };   // FuzzedKickoffState.h:7-27
class RandomState : public StateSetter { public:        // RandomState.h:5-16
    bool randBallSpeed, randCarSpeed, carsOnGround;
    RandomState(bool randBallSpeed, bool randCarSpeed, bool carsOnGround)
        : randBallSpeed(randBallSpeed), randCarSpeed(randCarSpeed), carsOnGround(carsOnGround) {}
    int DescribeKind() const override { return GGL_SETTER_RANDOM; }
    int DescribeFlags() const override {
        return (randBallSpeed ? GGL_RANDOM_BALL_SPEED : 0) | (randCarSpeed ? GGL_RANDOM_CAR_SPEED : 0) | (carsOnGround ? GGL_RANDOM_CARS_ON_GROUND : 0); } };
class CombinedState : public StateSetter { public:      // CombinedState.h:10-46 (children are not owned, as in the reference)
    std::vector<std::pair<StateSetter*, float>> setters;
    CombinedState(const std::vector<std::pair<StateSetter*, float>>& setters) : setters(setters) {
        if (setters.empty()) _Fatal("CombinedState: no setters");
        for (auto& p : setters)
            if (p.second < 0 || std::isnan(p.second) || std::isinf(p.second)) _Fatal("CombinedState: a state setter has a negative or invalid weight");
    }
    int DescribeKind() const override { return GGL_SETTER_COMBINED; } };

// ---------------------------------------------------------------------------------------------- EnvSet
struct EnvCreateResult {                                // EnvSet.h:14-23
    Arena* arena = nullptr;
    std::vector<WeightedReward> rewards;
    std::vector<TerminalCondition*> terminalConditions;
    ObsBuilder* obsBuilder = nullptr;
    ActionParser* actionParser = nullptr;
    StateSetter* stateSetter = nullptr;
    void* userInfo = nullptr;
};
typedef std::function<EnvCreateResult(int index)> EnvCreateFn;

struct EnvSetConfig {                                   // EnvSet.h:26-33
    EnvCreateFn envCreateFn;
    int numArenas = 0;
    int tickSkip = 8;
    int actionDelay = 7;
    bool saveRewards = false;
    bool shuffleRewardSampling = true;
    int device = 0;                                     // CUDA device ordinal (one EnvSet per GPU: arenas shard across ranks)
    uint64_t seed = 0;                                  // device RNG for kickoff permutations / respawn slots
};

struct EnvState {                                       // EnvSet.h:35-60 (host mirrors; GameState structs are fetched lazily)
    int numPlayers = 0;
    DimList2<float> obs;
    DimList2<uint8_t> actionMasks;
    std::vector<float> rewards;
    std::vector<std::vector<float>> lastRewards;
    std::vector<uint8_t> terminals;
    std::vector<int> arenaPlayerStartIdx;
};

// Translation of one EnvCreateResult into the device descriptor. Throws on anything the device cannot run.
inline ggl_envset_desc DescribeEnv(const EnvCreateResult& r, const EnvSetConfig& cfg) {
    ggl_envset_desc d;
    std::memset(&d, 0, sizeof(d));
    if (!r.arena) _Fatal("EnvCreateResult.arena is null");
    if (r.arena->gameMode != GameMode::SOCCAR && r.arena->gameMode != GameMode::HEATSEEKER)
        _Fatal("only GameMode::SOCCAR and GameMode::HEATSEEKER have a device integrator");
    d.game_mode = r.arena->gameMode == GameMode::HEATSEEKER ? GGL_MODE_HEATSEEKER : GGL_MODE_SOCCAR;
    int blue = 0, orange = 0; bool blueFirst = true;
    for (auto& c : r.arena->_cars) {
        if (c.team == Team::BLUE) { blue++; if (orange) blueFirst = false; } else orange++;
    }
    if (blue != orange || blue < 1 || blue > 3 || !blueFirst)
        _Fatal("the device EnvSet needs N blue cars added first, then N orange cars, N in 1..3");
    d.device = cfg.device; d.num_arenas = cfg.numArenas; d.players_per_team = blue;
    d.tick_skip = cfg.tickSkip; d.action_delay = cfg.actionDelay; d.seed = cfg.seed;
    if (!r.obsBuilder || r.obsBuilder->DescribeKind() < 0) _Fatal("ObsBuilder subclass has no device implementation (stock: DefaultObs, AdvancedObs, DefaultObsPadded)");
    d.obs_kind = r.obsBuilder->DescribeKind();
    d.obs_max_players = r.obsBuilder->DescribeMaxPlayers();
    r.obsBuilder->DescribeCoefs(d.obs_coefs);
    if (d.obs_kind == GGL_OBS_DEFAULT_PADDED && blue > d.obs_max_players)        // DefaultObsPadded.cpp:40-44
        _Fatal("DefaultObsPadded: too many players per team for maxPlayers");
    if (!r.actionParser || !r.actionParser->IsDefaultAction()) _Fatal("ActionParser subclass has no device implementation (stock: DefaultAction)");
    if (!r.stateSetter || r.stateSetter->DescribeKind() < 0) _Fatal("StateSetter subclass has no device implementation (stock: KickoffState, FuzzedKickoffState, RandomState, CombinedState; data: InjectedState)");
    d.state_setter = r.stateSetter->DescribeKind();
    d.state_setter_flags = r.stateSetter->DescribeFlags();
    if (d.state_setter == GGL_SETTER_COMBINED) {
        auto* comb = static_cast<const CombinedState*>(r.stateSetter);
        if (comb->setters.size() > GGL_MAX_SETTER_CHILDREN) _Fatal("CombinedState: more children than GGL_MAX_SETTER_CHILDREN");
        d.num_setter_children = (int)comb->setters.size();
        for (size_t i = 0; i < comb->setters.size(); i++) {
            const StateSetter* ch = comb->setters[i].first;
            if (!ch || ch->DescribeKind() < 0 || ch->DescribeKind() == GGL_SETTER_COMBINED)
                _Fatal("CombinedState: child has no device implementation (stock: KickoffState, FuzzedKickoffState, RandomState; data: InjectedState)");
            d.setter_children[i].kind = ch->DescribeKind(); d.setter_children[i].flags = ch->DescribeFlags();
            d.setter_children[i].weight = comb->setters[i].second;
        }
    }
    if (r.rewards.size() > GGL_MAX_REWARDS) _Fatal("too many reward functions");
    d.num_rewards = (int)r.rewards.size();
    for (size_t i = 0; i < r.rewards.size(); i++) {
        ggl_reward_desc& rd = d.rewards[i];
        if (!r.rewards[i].reward || !r.rewards[i].reward->Describe(rd))
            _Fatal("Reward subclass '" + (r.rewards[i].reward ? r.rewards[i].reward->GetName() : std::string("null")) + "' has no device implementation");
        rd.weight = r.rewards[i].weight;
    }
    if (r.terminalConditions.size() > GGL_MAX_TERMINALS) _Fatal("too many terminal conditions");
    d.num_terminals = (int)r.terminalConditions.size();
    for (size_t i = 0; i < r.terminalConditions.size(); i++)
        if (!r.terminalConditions[i] || !r.terminalConditions[i]->Describe(d.terminals[i]))
            _Fatal("TerminalCondition subclass has no device implementation (stock: GoalScoreCondition, NoTouchCondition)");
    return d;
}

// everything of a descriptor except pointers and the arena count must agree between arenas
inline bool SameDescriptor(ggl_envset_desc a, ggl_envset_desc b) {
    a.meshes = b.meshes = nullptr; a.num_meshes = b.num_meshes = 0;
    return std::memcmp(&a, &b, sizeof(a)) == 0;
}

struct EnvSet {                                         // EnvSet.h:62-110
    std::vector<Arena*> arenas;
    std::vector<void*> userInfos;
    EnvSetConfig config;
    int obsSize = 0, numActions = 0;
    std::vector<std::vector<WeightedReward>> rewards;
    std::vector<std::vector<TerminalCondition*>> terminalConditions;
    std::vector<ObsBuilder*> obsBuilders;
    std::vector<ActionParser*> actionParsers;
    std::vector<StateSetter*> stateSetters;
    EnvState state;
    ggl_envset_desc desc;

    explicit EnvSet(const EnvSetConfig& cfg) : config(cfg) {
        if (cfg.numArenas <= 0 || !cfg.envCreateFn) _Fatal("EnvSetConfig: numArenas and envCreateFn are required");
        for (int i = 0; i < cfg.numArenas; i++) {       // EnvSet.cpp:46-104: one envCreateFn call per arena
            EnvCreateResult r = cfg.envCreateFn(i);
            ggl_envset_desc di = DescribeEnv(r, cfg);
            if (i == 0) desc = di;
            else if (!SameDescriptor(desc, di)) _Fatal("arena " + std::to_string(i) + " is configured differently from arena 0 (the device EnvSet is homogeneous)");
            arenas.push_back(r.arena); userInfos.push_back(r.userInfo); rewards.push_back(r.rewards);
            terminalConditions.push_back(r.terminalConditions); obsBuilders.push_back(r.obsBuilder);
            actionParsers.push_back(r.actionParser); stateSetters.push_back(r.stateSetter);
        }
        auto& meshes = RocketSim::_Meshes();
        for (auto& m : meshes) _blobs.push_back(ggl_mesh_blob{m.data(), m.size()});
        desc.num_meshes = (int)_blobs.size(); desc.meshes = _blobs.data();
        _Check(ggl_envset_create(&desc, &_h), "ggl_envset_create");
        _Check(ggl_envset_views_get(_h, &_views), "ggl_envset_views_get");
        obsSize = _views.obs_size; numActions = _views.num_actions;
        state.numPlayers = _views.num_players;
        for (int i = 0; i < cfg.numArenas; i++) state.arenaPlayerStartIdx.push_back(i * _views.players_per_arena);
        state.obs.Resize(state.numPlayers, obsSize);
        state.actionMasks.Resize(state.numPlayers, numActions);
        state.rewards.assign(state.numPlayers, 0.f);
        state.terminals.assign(cfg.numArenas, 0);
        state.lastRewards.assign(cfg.numArenas, std::vector<float>(desc.num_rewards, 0.f));
        _actions.assign(state.numPlayers, 0);
        _Pull(true);                                     // the constructor's Reset built obs + masks (EnvSet.cpp:106-110)
    }
    EnvSet(const EnvSet&) = delete;
    EnvSet& operator=(const EnvSet&) = delete;
    ~EnvSet() {
        if (_h) ggl_envset_destroy(_h);
        for (Arena* a : arenas) delete a;               // EnvSet owns arenas, never the plugins (EnvSet.h:93-101)
    }

    void StepFirstHalf(bool async) {                    // EnvSet.cpp:113-130
        _Check(ggl_envset_step_first_half(_h, nullptr), "ggl_envset_step_first_half");
        if (!async) Sync();
    }
    void StepSecondHalf(const IList& actionIndices, bool async) {   // EnvSet.cpp:132-273
        if ((int)actionIndices.size() != state.numPlayers) _Fatal("StepSecondHalf: actionIndices.size() != numPlayers");
        for (int a : actionIndices)
            if (a < 0 || a >= numActions) _Fatal("StepSecondHalf: action index out of range");
        _Check(ggl_envset_step_second_half_host(_h, actionIndices.data()), "ggl_envset_step_second_half_host");
        _Pull(false);
        (void)async;                                    // host mirrors are complete on return either way
    }
    void Sync() { _Check(ggl_envset_sync(_h), "ggl_envset_sync"); }
    void Reset() {                                      // EnvSet.cpp:338-354: arenas whose terminal != 0
        _Check(ggl_envset_reset(_h, nullptr), "ggl_envset_reset");
        _Pull(true);
    }
    void ResetArena(int index) {                        // EnvSet.cpp:275-331 called directly: unconditional
        _Check(ggl_envset_reset_arena(_h, index, nullptr), "ggl_envset_reset_arena");
        _Pull(true);
    }
    // device buffers (what torch::from_blob wraps zero-copy), valid until the EnvSet is destroyed
    const ggl_envset_views& deviceViews() const { return _views; }
    ggl_envset* handle() const { return _h; }
    // full physics state of one arena (StateSetter output / stepCallback input), see ggl_arena_state
    void GetArenaState(int arena, ggl_arena_state& out) { _Check(ggl_envset_get_state(_h, arena, &out), "ggl_envset_get_state"); }
    void SetArenaState(int arena, const ggl_arena_state& in) { _Check(ggl_envset_set_state(_h, arena, &in), "ggl_envset_set_state"); }
    void SetResetState(int arena, const ggl_arena_state& in) { _Check(ggl_envset_set_reset_state(_h, arena, &in), "ggl_envset_set_reset_state"); }

private:
    ggl_envset* _h = nullptr;
    ggl_envset_views _views;
    std::vector<ggl_mesh_blob> _blobs;
    std::vector<int32_t> _actions;
    static void _Check(int rc, const char* what) {
        if (rc != GGL_OK) _Fatal(std::string(what) + " -> " + ggl_last_error());
    }
    void _Pull(bool afterReset) {
        _Check(ggl_envset_read_host(_h, state.obs.data.data(), state.actionMasks.data.data(), afterReset ? nullptr : state.rewards.data(),
                                    state.terminals.data()), "ggl_envset_read_host");
    }
};

}  // namespace RLGC
