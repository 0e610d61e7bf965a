"""World-model training entry point: same command line, model selection, loss wiring and output files as the reference's
/root/reference/train_env_model.py:20-130 (`--model baseline | multistep | termination_predictor`), with the trainers
running on the device (envmodel/*_trainer.py -> libfqlpop.so). One model more than the reference wires:
`--model initial_observation` trains the VAE of envmodel/initial_observation.py on the episode-start observations
(InitialObservationLoader) with the flags the reference's parser already has for it (--model.latent_dim,
--reconstruction_weight) and leaves initial_observation.pt / _config.yaml where distribution_visualizer.py:61-94 reads
them. Differences: datasets come from
fql.envs.env_utils.make_env_and_datasets (ogbench files if present in --data_directory, synthetic otherwise; ogbench itself
is not installed here), and wandb is optional (`--use_wandb`)."""
from dataclasses import asdict
from functools import partial

from argparser import build_env_model_config_from_args, get_env_model_argparser
from envmodel.baseline import BaselineStatePredictor, TerminationPredictor
from envmodel.initial_observation import InitialObservationEnvModel, InitialObservationTrainer, vae_loss
from envmodel.loss import focal_loss, state_prediction_loss
from envmodel.multistep import MultistepStatePredictor
from envmodel.state_predictor_trainer import StatePredictorTrainer
from envmodel.termination_predictor_trainer import TerminationPredictorTrainer
from fql.envs.env_utils import make_env_and_datasets
from utils.data_loader import InitialObservationLoader, MultistepLoader, StepLoader
from utils.envmodel import save_model


def main(argv=None):
    parser = get_env_model_argparser()
    parser.add_argument("--use_wandb", action="store_true", help="Log to Weights & Biases (if installed).")
    parser.add_argument("--device_sampling", action="store_true",
                        help="Keep the training set in HBM and draw batches (or multistep windows) on the device.")
    args = parser.parse_args(argv)
    use_wandb, device_sampling = args.use_wandb, args.device_sampling
    del args.use_wandb, args.device_sampling
    config = build_env_model_config_from_args(args)
    made = make_env_and_datasets(config.env_name, config.data_directory)
    train_dataset, val_dataset = made[-2], made[-1]
    state_loss = partial(state_prediction_loss, true_termination_weight=config.true_termination_weight,
                         termination_weight=config.termination_weight, reconstruction_weight=0.0)
    if config.model == "baseline":
        train_loader, val_loader = StepLoader(train_dataset), StepLoader(val_dataset)
        sample = train_loader.sample(config.batch_size)
        model = BaselineStatePredictor(sample["observations"].shape[-1], sample["actions"].shape[-1],
                                       tuple(config.model_config["hidden_dims"]))
        loss_fn = state_loss
    elif config.model == "multistep":
        train_loader = MultistepLoader(train_dataset, sequence_length=config.sequence_length)
        val_loader = MultistepLoader(val_dataset, sequence_length=config.sequence_length)
        sample = train_loader.sample(config.batch_size)
        model = MultistepStatePredictor(sample["observations"].shape[-1], sample["actions"].shape[-1],
                                        tuple(config.model_config["hidden_dims"]))
        loss_fn = state_loss
    elif config.model == "termination_predictor":
        train_loader, val_loader = StepLoader(train_dataset), StepLoader(val_dataset)
        sample = train_loader.sample(config.batch_size)
        model = TerminationPredictor(sample["observations"].shape[-1], tuple(config.model_config["hidden_dims"]))
        loss_fn = focal_loss
    elif config.model == "initial_observation":
        train_loader, val_loader = InitialObservationLoader(train_dataset), InitialObservationLoader(val_dataset)
        sample = train_loader.sample(config.batch_size)
        model = InitialObservationEnvModel(sample["observations"].shape[-1], int(config.model_config["latent_dim"]),
                                           tuple(config.model_config["hidden_dims"]))
        loss_fn = partial(vae_loss, reconstruction_weight=config.reconstruction_weight)
    else:
        raise ValueError(f"Unknown model type: {config.model}")
    logger = None
    if use_wandb:
        import wandb
        logger = wandb.init(name=f"{config.env_name}_{config.model}", project="fql_env_model", config=asdict(config), dir="exp/")
    cls = {"termination_predictor": TerminationPredictorTrainer,
           "initial_observation": InitialObservationTrainer}.get(config.model, StatePredictorTrainer)
    trainer = cls(model=model, train_loader=train_loader, val_loader=val_loader, loss_fn=loss_fn, config=config,
                  logger=logger, device_sampling=device_sampling)
    trainer.train()
    path = save_model(trainer.state.params, config.save_directory, config.env_name, config.model, config.model_config)
    print(f"saved {path}; validation: {trainer.last_validation}")
    if logger is not None:
        logger.finish()
    return trainer


if __name__ == "__main__":
    main()
