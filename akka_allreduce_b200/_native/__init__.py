"""Built native libraries live here (``*.so`` are git-ignored; see ``akka_allreduce_b200.build``)."""
