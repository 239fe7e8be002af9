"""`python main_ddp.py ...` -- same command line as the reference's main_ddp.py; the work happens in
distributed_pytorch_b200.cli.run("ddp")."""
from distributed_pytorch_b200.cli import main

if __name__ == "__main__":
    main("ddp")
