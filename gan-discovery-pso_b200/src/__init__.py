"""Drop-in mirror of the reference's `src` package for the swarm-evaluation hot path:
`src.pso.util_pso` (Particle, Swarm), `src.pso.util_discovery` (Discovery),
`src.utils.util_dcgan` / `src.models` (Generator, Discriminator), `src.hands_on.pso`.
Put `gan-discovery-pso_b200/` on sys.path in place of the reference root."""
