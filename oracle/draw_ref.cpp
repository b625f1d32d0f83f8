// ORACLE (test infrastructure, not product code) -- pin for the multinomial draw.
//
// llama.cpp's llama_sample_token (reached from create_chat_completion at /root/reference/api.py:55-63
// through llama-cpp-python 0.2.77; un-vendored upstream) ends with
//     std::discrete_distribution<> dist(probs.begin(), probs.end());  int idx = dist(rng);
// on the context's std::mt19937. This program IS that call, made with the same libstdc++ <random> the
// reference build links, so its output pins oracle/sampler_ref.py's `draw` and `MT19937Uniforms`
// (tests/golden/make_sampler_golden.py writes tests/golden/sampler_draws.json from it;
// tests/test_oracle_sampler.py compares, and re-runs this program when g++ is available).
//
// stdin:  n_cases, then per case: seed n_draws n_probs p0 p1 ... (probabilities as float32 text)
// stdout: per case one line of n_draws indices; one mt19937 per case, reused across its draws, and a
//         fresh distribution object per draw (llama.cpp constructs one per sampled token).
#include <cstdio>
#include <random>
#include <vector>

int main() {
    int n_cases = 0;
    if (scanf("%d", &n_cases) != 1) return 1;
    for (int c = 0; c < n_cases; c++) {
        unsigned seed; int n_draws, n;
        if (scanf("%u %d %d", &seed, &n_draws, &n) != 3) return 1;
        std::vector<float> probs(n);
        for (int i = 0; i < n; i++) if (scanf("%f", &probs[i]) != 1) return 1;
        std::mt19937 rng(seed);
        for (int d = 0; d < n_draws; d++) {
            std::discrete_distribution<> dist(probs.begin(), probs.end());
            printf("%d%c", dist(rng), d + 1 == n_draws ? '\n' : ' ');
        }
    }
    return 0;
}
