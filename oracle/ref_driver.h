/* oracle/ref_driver.h — POD config shared by oracle/ref_driver.cpp and its ctypes mirror
 * (tests/oracle_ref.py).  TEST INFRASTRUCTURE ONLY. */
#ifndef CPPRL_ORACLE_REF_DRIVER_H
#define CPPRL_ORACLE_REF_DRIVER_H
#include <stdint.h>

enum { REF_ALGO_PPO = 0, REF_ALGO_A2C = 1 };
enum { REF_BASE_MLP = 0, REF_BASE_GRU_MLP = 1, REF_BASE_CNN = 2 };
enum { REF_HEAD_CATEGORICAL = 0, REF_HEAD_NORMAL = 1 };

typedef struct RefConfig
{
    int32_t algo, base, head;
    int32_t T, N;
    int32_t obs_ndim;
    int64_t obs_shape[3];   /* {D} for MLP bases, {C,H,W} for CNN */
    int64_t num_actions;    /* Discrete: number of choices; Box: action dims */
    int32_t hidden;         /* policy hidden size (64 / 512) */
    int32_t storage_hidden; /* RolloutStorage hidden_state_size */
    int32_t normalize_obs;
    int32_t num_epoch, num_mini_batch;
    float clip_param, actor_loss_coef, value_loss_coef, entropy_coef;
    float learning_rate, epsilon, alpha, max_grad_norm, kl_target;
} RefConfig;

#endif
